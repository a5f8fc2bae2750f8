"""Reference-shaped model modules (reference: models/vit.py, models/vit_no_classifier.py)."""
