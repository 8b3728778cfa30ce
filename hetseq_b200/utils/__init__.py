"""Utilities: meters, progress/log sinks, checkpoint I/O, misc helpers."""
