"""Empty stand-in (import-only); see oracle/refshim/README.md."""
