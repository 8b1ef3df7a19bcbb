"""example chemistries (python/magicsoup/examples of the reference): data for tests and benchmarks"""
