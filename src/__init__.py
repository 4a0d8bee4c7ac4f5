"""Reference-compatible import paths (``from src.trainer import Trainer`` ...), backed by ml_trainer_b200."""
