"""Placeholder module, empty in the reference too (sr/langmodel/langmodel.py)."""
__all__ = []
