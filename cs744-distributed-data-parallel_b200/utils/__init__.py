"""Support code: log lines, timers, seeding, checkpoints."""
