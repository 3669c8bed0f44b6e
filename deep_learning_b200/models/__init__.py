"""Reference-shaped model modules: `import deep_learning_b200.models.<alg_name> as alg_model`
mirrors `import models.<alg_name> as alg_model` (local_run_textline.py:142-165)."""
