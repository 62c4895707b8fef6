"""reportree_b200 -- B200-native allele-matrix partitioning hot path."""
__version__ = "0.1.0"
