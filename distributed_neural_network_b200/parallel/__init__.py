"""Process bootstrap, peer-memory group, membership epochs, launcher."""
