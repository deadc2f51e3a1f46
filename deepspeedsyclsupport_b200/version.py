__version__ = "0.12.7+b200.1"
__version_major__, __version_minor__, __version_patch__ = 0, 12, 7
