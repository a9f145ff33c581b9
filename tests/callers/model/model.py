raise ImportError("tests/callers/model/model.py was imported: the script directory shadowed the zero_tig_b200 "
                  "drop-in (use `python -m zero_tig_b200.run <script>`)")
