"""TEST INFRASTRUCTURE ONLY — empty matplotlib.pyplot stub."""
