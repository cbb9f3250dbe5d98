"""Estimator layer: config, checkpoints (TF variable names), hooks, train/evaluate/predict, best-model export."""
