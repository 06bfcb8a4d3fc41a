from ..managers import ClientManager  # reference import path

__all__ = ["ClientManager"]
