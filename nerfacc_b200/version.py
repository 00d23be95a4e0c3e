"""Version of the B200-native hot path; tracks the reference API level it mirrors (nerfacc 0.5.3)."""
__version__ = "0.5.3+b200.0"
