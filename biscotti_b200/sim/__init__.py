"""ML-only sandboxes (in-process training loops without ledger or crypto)."""
