"""HF model integration: same entry points as the reference's models/ package."""
