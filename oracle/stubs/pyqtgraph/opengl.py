"""Import stub for `import pyqtgraph.opengl as gl` (plot.py:3)."""
