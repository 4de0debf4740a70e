"""tksm_b200 — B200-native implementation of the `tksm sequence` (Seq exit module) hot path."""
import os

MODELS_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "models")
