"""tachikoma_b200: B200-native drop-in for the tachikoma BYOC hot path (QNN int8 operator chain)."""
from . import relay, tachikoma  # noqa: F401
from .runtime import TachikomaJSONRuntime, TachikomaError, create_executor  # noqa: F401
