"""Per-process runtime: vertex operator libraries (host + GPU), executor, backends, host library bindings."""
