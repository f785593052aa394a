from .arena import ParamArena, BufferArena

__all__ = ["ParamArena", "BufferArena"]
