"""B200-native screen fitting: h5parm solutions -> FITS a-term cubes on hand-written CUDA."""
__version__ = "0.1.0"
