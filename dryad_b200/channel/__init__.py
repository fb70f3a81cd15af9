"""Channels and storage: record codec, block streams, partfile tables, text files, FIFOs."""
