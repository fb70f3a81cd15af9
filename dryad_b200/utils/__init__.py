"""Small shared utilities (hashing, logging, config, bit vectors)."""
