"""See tests/stubs/matplotlib/__init__.py."""
