"""URDF -> ExoGeometry. Reads the kinematic tree the way pybullet's importer numbers it
(reference: Exoskeleton_sim_pybullet.py:14-18 loads Simulation/exo_v3.urdf; :21-25,48-63 hard-code the
resulting link indices). Only what the env's position reads need is kept: joint origins, axes, limits and
inertial (COM) offsets, since getLinkState(...)[0] returns the COM position."""
import json
import math
import os
from xml.dom import minidom

from . import _capi

# Exoskeleton_sim_pybullet.py:48-63 — handles in the order act11, act12, act21, act22, ..., act71, act72
ACT_LINKS = (9, 5, 12, 6, 15, 8, 17, 11, 14, 7, 18, 13, 16, 10)
SHOULDER_LINK, ELBOW_LINK = 0, 3


def _floats(text, n=3, default=0.0):
    if not text:
        return [default] * n
    return [float(t) for t in text.split()]


def _rotation_from_rpy(roll, pitch, yaw):
    sr, cr = math.sin(roll), math.cos(roll)
    sp, cp = math.sin(pitch), math.cos(pitch)
    sy, cy = math.sin(yaw), math.cos(yaw)
    # fixed-axis X-Y-Z == Rz(yaw) Ry(pitch) Rx(roll), written out
    return [cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr,
            sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr,
            -sp, cp * sr, cp * cr]


def _child(elem, tag):
    for c in elem.childNodes:
        if c.nodeType == c.ELEMENT_NODE and c.tagName == tag:
            return c
    return None


def parse_urdf(path, base_position=(0.0, 0.0, 0.1)):
    """Returns a plain dict (JSON-serialisable) describing the tree in Bullet link order."""
    doc = minidom.parse(path)
    robot = doc.documentElement
    com, dyn = {}, {}
    for ln in robot.getElementsByTagName("link"):
        inertial = _child(ln, "inertial")
        org = _child(inertial, "origin") if inertial is not None else None
        com[ln.getAttribute("name")] = _floats(org.getAttribute("xyz")) if org is not None else [0.0, 0.0, 0.0]
        # inertial data for the general multibody step (EXO_MOTOR_FULL): mass, principal inertias, inertial-frame rotation,
        # and the collision box pybullet derives its default inertia from
        rec = dict(mass=0.0, inertia_diag=[0.0] * 3, inertial_R=_rotation_from_rpy(0.0, 0.0, 0.0),
                   box_size=[0.0] * 3, box_xyz=[0.0] * 3, box_R=_rotation_from_rpy(0.0, 0.0, 0.0))
        if inertial is not None:
            ms, it = _child(inertial, "mass"), _child(inertial, "inertia")
            rec["mass"] = float(ms.getAttribute("value")) if ms is not None else 0.0
        if inertial is not None and _child(inertial, "inertia") is not None:
            it = _child(inertial, "inertia")
            if any(float(it.getAttribute(k) or 0.0) != 0.0 for k in ("ixy", "ixz", "iyz")):
                raise ValueError("link %s: products of inertia are not supported" % ln.getAttribute("name"))
            rec["inertia_diag"] = [float(it.getAttribute(k)) for k in ("ixx", "iyy", "izz")]
            if org is not None and org.hasAttribute("rpy"):
                rec["inertial_R"] = _rotation_from_rpy(*_floats(org.getAttribute("rpy")))
        col = _child(ln, "collision")
        if col is not None:
            geom = _child(col, "geometry")
            box = _child(geom, "box") if geom is not None else None
            corg = _child(col, "origin")
            if box is not None:
                rec["box_size"] = _floats(box.getAttribute("size"))
            if corg is not None:
                rec["box_xyz"] = _floats(corg.getAttribute("xyz"))
                rec["box_R"] = _rotation_from_rpy(*_floats(corg.getAttribute("rpy")))
        dyn[ln.getAttribute("name")] = rec
    by_parent, children = {}, set()
    for jn in robot.getElementsByTagName("joint"):
        parent = _child(jn, "parent").getAttribute("link")
        child = _child(jn, "child").getAttribute("link")
        org, axis, lim = _child(jn, "origin"), _child(jn, "axis"), _child(jn, "limit")
        rec = dict(name=jn.getAttribute("name"), kind=jn.getAttribute("type"), child=child,
                   xyz=_floats(org.getAttribute("xyz")) if org is not None else [0.0] * 3,
                   rpy=_floats(org.getAttribute("rpy")) if org is not None else [0.0] * 3,
                   axis=_floats(axis.getAttribute("xyz")) if axis is not None else [1.0, 0.0, 0.0],
                   lower=float(lim.getAttribute("lower")) if lim is not None and lim.hasAttribute("lower") else -1e30,
                   upper=float(lim.getAttribute("upper")) if lim is not None and lim.hasAttribute("upper") else 1e30)
        by_parent.setdefault(parent, []).append(rec)
        children.add(child)
    roots = [name for name in com if name not in children]
    if len(roots) != 1:
        raise ValueError("URDF must have exactly one root link, found %r" % roots)
    order = []          # pre-order DFS, children in file order: Bullet's multibody link numbering
    stack = [(rec, -1) for rec in reversed(by_parent.get(roots[0], []))]
    while stack:
        rec, parent_idx = stack.pop()
        idx = len(order)
        order.append((rec, parent_idx))
        for sub in reversed(by_parent.get(rec["child"], [])):
            stack.append((sub, idx))
    kinds = {"revolute": 0, "continuous": 0, "prismatic": 1}
    links = []
    for rec, parent_idx in order:
        if rec["kind"] not in kinds:
            raise ValueError("joint %s: type %s is not supported" % (rec["name"], rec["kind"]))
        links.append(dict(name=rec["child"], joint=rec["name"], parent=parent_idx, joint_type=kinds[rec["kind"]],
                          origin_xyz=rec["xyz"], origin_R=_rotation_from_rpy(*rec["rpy"]), axis=rec["axis"],
                          com_xyz=com[rec["child"]], lower=rec["lower"], upper=rec["upper"], **dyn[rec["child"]]))
    base_com = com[roots[0]]
    return dict(base=roots[0], base_pos=[base_position[i] - base_com[i] for i in range(3)], links=links,
                act_link=list(ACT_LINKS), shoulder_link=SHOULDER_LINK, elbow_link=ELBOW_LINK)


def to_struct(tree):
    if len(tree["links"]) != _capi.NLINKS:
        raise ValueError("expected %d links (exo_v3.urdf), got %d" % (_capi.NLINKS, len(tree["links"])))
    g = _capi.ExoGeometry()
    for i, ln in enumerate(tree["links"]):
        g.parent[i] = ln["parent"]
        g.joint_type[i] = ln["joint_type"]
        g.lower[i], g.upper[i] = ln["lower"], ln["upper"]
        for k in range(3):
            g.origin_xyz[i][k] = ln["origin_xyz"][k]
            g.axis[i][k] = ln["axis"][k]
            g.com_xyz[i][k] = ln["com_xyz"][k]
        for k in range(9):
            g.origin_R[i][k] = ln["origin_R"][k]
    for k in range(3):
        g.base_pos[k] = tree["base_pos"][k]
    for a in range(14):
        g.act_link[a] = tree["act_link"][a]
    g.shoulder_link, g.elbow_link = tree["shoulder_link"], tree["elbow_link"]
    return g


def to_dynamics_struct(tree):
    """ExoLinkDynamics (include/exo_env.h) for exo_set_motor_model."""
    d = _capi.ExoLinkDynamics()
    for i, ln in enumerate(tree["links"]):
        if "mass" not in ln:
            raise ValueError("geometry table has no inertial data (regenerate it with tools/extract_reference_data.py)")
        d.mass[i] = ln["mass"]
        for k in range(3):
            d.inertia_diag[i][k] = ln["inertia_diag"][k]
            d.box_size[i][k] = ln["box_size"][k]
            d.box_xyz[i][k] = ln["box_xyz"][k]
        for k in range(9):
            d.inertial_R[i][k] = ln["inertial_R"][k]
            d.box_R[i][k] = ln["box_R"][k]
    return d


BUNDLED = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "exo_v3_geometry.json")


def load_geometry(urdf_path=None):
    """Geometry from a URDF file, or (default) from the bundled table that tools/extract_reference_data.py
    derived from the reference's Simulation/exo_v3.urdf with parse_urdf()."""
    if urdf_path is not None:
        return parse_urdf(urdf_path)
    with open(BUNDLED) as f:
        return json.load(f)
