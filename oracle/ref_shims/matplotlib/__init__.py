"""Shim: imported at ragen/env/sokoban/utils.py:8-9, used only by plotting helpers."""
