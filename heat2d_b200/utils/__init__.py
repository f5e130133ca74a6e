"""Dump writers, timing and clock sampling helpers."""
