from .cache import Cache
