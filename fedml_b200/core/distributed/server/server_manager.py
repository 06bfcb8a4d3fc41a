from ..managers import ServerManager  # reference import path

__all__ = ["ServerManager"]
