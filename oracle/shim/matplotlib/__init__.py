"""ORACLE shim: empty stand-in so that `import rafft` (rafft/__init__.py:2 ->
rafft/rafft_kin.py:14 `import matplotlib.pyplot`) succeeds without matplotlib."""
