"""TF1-format (tensor bundle) checkpoint reader / writer."""
from .bundle import BundleReader, BundleWriter, merge_bundles  # noqa: F401
