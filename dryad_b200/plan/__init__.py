"""Query compiler (L1): logical trees, dataset properties, rewrites, decomposition, physical plan."""
