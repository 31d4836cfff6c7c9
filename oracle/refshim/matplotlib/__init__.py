"""TEST INFRASTRUCTURE ONLY — empty stand-in for matplotlib (render() is out of scope)."""
cm = None
colormaps = {}
