"""Timing, clock sampling."""
