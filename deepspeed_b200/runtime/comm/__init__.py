"""Communication algorithms used by optimizers (reference ``runtime/comm``): 1-bit compressed all-reduce, coalesced
reduce-scatter."""
