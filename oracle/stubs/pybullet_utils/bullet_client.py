"""Shim for `pybullet_utils.bullet_client.BulletClient` (reference call sites:
Environment/Exoskeleton_env.py:113-118,433,484 and
Environment/Exoskeleton_sim_pybullet.py:18,115-117,129-142,194-195,348-349).
Every call is forwarded to oracle/minibullet.py."""
import os
import sys

_here = os.path.dirname(os.path.abspath(__file__))
_oracle = os.path.abspath(os.path.join(_here, "..", ".."))
if _oracle not in sys.path:
    sys.path.insert(0, _oracle)
import minibullet  # noqa: E402

# knobs for the restated Bullet boundary; gen_golden.py may override before constructing envs
DEFAULT_KNOBS = dict(link_state_lag=1, joint_limits="off")


class BulletClient(object):
    _next_id = 0

    def __init__(self, connection_mode=None, options=""):
        if connection_mode == 1:
            raise RuntimeError("p.GUI needs a graphics stack: unavailable on this box")
        self._client = BulletClient._next_id
        BulletClient._next_id += 1
        self._world = minibullet.World(**DEFAULT_KNOBS)

    def setAdditionalSearchPath(self, path):
        pass

    def resetSimulation(self):
        self._world.reset_simulation()

    def setGravity(self, gx, gy, gz):
        self._world.gravity = (gx, gy, gz)

    def setTimeStep(self, dt):
        self._world.dt = float(dt)

    def setRealTimeSimulation(self, flag):
        pass

    def loadURDF(self, fileName, basePosition=(0, 0, 0), physicsClientId=0, useFixedBase=False, **kw):
        if not useFixedBase:
            raise NotImplementedError("minibullet models the fixed-base chain only")
        return self._world.load_urdf(fileName, basePosition)

    def setJointMotorControlArray(self, bodyUniqueId, jointIndices, controlMode=None, targetPositions=None, **kw):
        if controlMode != 2 or kw:
            raise NotImplementedError("only POSITION_CONTROL with default gains is restated")
        self._world.set_position_targets(bodyUniqueId, list(jointIndices), [float(x) for x in targetPositions])

    def getLinkState(self, bodyUniqueId, linkIndex=None, **kw):
        if kw.get("computeForwardKinematics", 0):
            raise NotImplementedError
        pos = self._world.link_com_world(bodyUniqueId, linkIndex)
        return (tuple(float(x) for x in pos),)

    def stepSimulation(self):
        self._world.step_simulation()

    def disconnect(self):
        pass
