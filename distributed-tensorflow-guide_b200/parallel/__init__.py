"""Parallel layer: transport, peer-memory fabric, local<->global helpers and the
scheme engines."""
