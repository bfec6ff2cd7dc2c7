"""Data modules."""
