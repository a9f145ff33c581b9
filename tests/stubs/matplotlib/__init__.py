"""Empty stand-in for matplotlib (absent from this image).  The reference's utils/utils.py:6 imports
``matplotlib.pyplot`` at module scope but only ``show_pic`` uses it; tests that run the unmodified reference
CLIs put tests/stubs on PYTHONPATH.  Test infrastructure only."""
