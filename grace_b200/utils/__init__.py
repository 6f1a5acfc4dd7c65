from grace_b200.utils.philox import fnv1a, philox_uniform, philox_normal, philox_u32, random_subset  # noqa: F401
