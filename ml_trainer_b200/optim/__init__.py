from .fused import FusedOptimizer
from .rules import RULE_IDS, Hyper, apply_rule

__all__ = ["FusedOptimizer", "RULE_IDS", "Hyper", "apply_rule"]
