"""``main.py``-style entry points, one per reference part."""
