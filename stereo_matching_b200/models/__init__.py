from .model import base_model, base_model_ws_9  # noqa: F401
