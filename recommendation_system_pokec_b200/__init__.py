"""B200-native FAS scoring path for pymlex/recommendation-system-pokec."""
from .world import World, FIXED_FIELDS, NUM_FIXED  # noqa: F401
