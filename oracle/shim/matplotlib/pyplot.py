"""ORACLE shim: plotting is unavailable on this box; importing must merely succeed."""
