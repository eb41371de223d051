"""Data-efficiency library (reference ``runtime/data_pipeline/``): curriculum learning (legacy + sampler-based),
random layerwise token dropping, dynamic batching with LR scaling, mmap indexed datasets, offline data analyzer."""
