"""urgym_b200 — B200-native batched physics backend for the URGym step path."""
from ._lib import ENV_TYPES, UrgymError  # noqa: F401
