"""Device ops: thin wrappers over the sm_100a kernels (see kernels.py) and the merge pipeline."""
