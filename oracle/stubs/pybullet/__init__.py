"""Shim for `import pybullet as p` (reference: Environment/Exoskeleton_env.py:3,
Environment/Exoskeleton_sim_pybullet.py:1). Only the constants the env touches."""
DIRECT = 2
GUI = 1
POSITION_CONTROL = 2
VELOCITY_CONTROL = 0
TORQUE_CONTROL = 1
