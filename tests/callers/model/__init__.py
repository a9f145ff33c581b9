"""Decoy: plays the role of the reference tree's own ``model`` package, which sits in the script directory and
therefore at ``sys.path[0]`` of ``python predict.py``.  Importing it means the drop-in did NOT win."""
