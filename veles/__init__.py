"""veles namespace: holds simd_b200, the B200-native convolve/correlate hot path."""
