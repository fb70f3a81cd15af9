"""Client API (L0): context, lazy queries, record types, decomposable protocols, error codes."""
