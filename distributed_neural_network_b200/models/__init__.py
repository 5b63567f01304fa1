from .network import Network  # noqa: F401
