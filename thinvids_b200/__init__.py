"""thinvids_b200 — B200-native scale / pixel-format stage for the thinvids encode task."""
