"""Training loop with the reference's API (reference: going_modular/going_modular/engine.py)."""
