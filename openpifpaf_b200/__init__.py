"""B200-native CIF/CAF pose decoder (drop-in for the openpifpaf CifCaf decoder)."""
__version__ = "0.1.0"
