"""Configuration layer (schema + loader)."""

from llmtrain.config import loader, schemas

__all__ = ["loader", "schemas"]
