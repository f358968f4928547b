"""Network blocks / models: the reference's transformer blocks (`replicann.arch.transformer`) and the
classic CNNs BASELINE.json names.  Import by full module path, as with the reference."""
