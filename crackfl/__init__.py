"""crackfl: Blackwell-native federated crack segmentation (see README.md / DESIGN.md)."""
__version__ = "0.1.0"
