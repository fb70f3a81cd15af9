"""Multi-GPU substrate: symmetric memory, fused shuffle, distributed sort / aggregate / join."""
