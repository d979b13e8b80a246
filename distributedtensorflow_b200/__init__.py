"""distributedtensorflow_b200 — a B200-native parameter-server training framework with the
capabilities of SvenGronauer/distributedTensorFlow (TF1 between-graph replication on MNIST).

See DESIGN.md for the architecture and SURVEY.md for the reference inventory.
"""
__version__ = "0.1.0"
