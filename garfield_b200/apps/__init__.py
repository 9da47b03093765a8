"""Application logic shared by the `applications/*/trainer.py` entry points."""
