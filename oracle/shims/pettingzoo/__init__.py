__version__ = "1.24.2"
