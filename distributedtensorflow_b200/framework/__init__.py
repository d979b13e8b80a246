"""Deferred-execution graph (TF1-style surface) and its backend executors."""
