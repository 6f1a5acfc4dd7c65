"""Data parallelism -- the one strategy GRACE has: bucketing, overlap with
backward, and the NVLink peer-memory engine that replaces the collective library."""
