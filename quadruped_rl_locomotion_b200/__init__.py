"""B200-native batched simulator for the Go1 locomotion environment (drop-in for the reference's
``Go1MujocoEnv`` + ``SubprocVecEnv`` hot path).  See DESIGN.md."""
from .mjcf import Model, compile_mjcf  # noqa: F401

__all__ = ["Model", "compile_mjcf"]
