"""rayca_b200 — B200-native backend for rayca-soft's per-pixel raytracing hot path."""
