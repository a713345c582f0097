"""Reference-motion tables. File format of Simulation/reference_motions/ref_motion_<n>.txt: one row per
1/40 s sample, 17 whitespace-separated columns; the env uses columns 0-4 = elbow_y, elbow_z, shoulder_x,
shoulder_y, shoulder_z in degrees (Utilities/read_txt_env.py:107-125; the acceleration columns 5-16 are parsed
there but never read by the env)."""
import os

import numpy as np

BUNDLED = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "reference_motions.npz")
ROWS = ("elbow_joint_y_positions", "elbow_joint_z_positions", "shoulder_joint_x_positions",
        "shoulder_joint_y_positions", "shoulder_joint_z_positions")


def read_motion_file(path):
    """-> float64 array [5, L] in the row order of ROWS."""
    cols = [[] for _ in range(5)]
    with open(path) as f:
        for line in f:
            parts = line.split()
            if not parts:
                continue
            if len(parts) < 17:
                raise ValueError("%s: expected 17 columns, got %d" % (path, len(parts)))
            for r in range(5):
                cols[r].append(float(parts[r]))
    return np.asarray(cols, dtype=np.float64)


def load_motions(source=None, count=8):
    """source: None -> the bundled tables; a directory -> ref_motion_0..count-1.txt inside it;
    a list of [5,L] arrays -> used as is."""
    if source is None:
        z = np.load(BUNDLED)
        return [np.ascontiguousarray(z["motion_%d" % m]) for m in range(len(z.files))]
    if isinstance(source, (str, os.PathLike)):
        return [read_motion_file(os.path.join(source, "ref_motion_%d.txt" % m)) for m in range(count)]
    return [np.ascontiguousarray(a, dtype=np.float64) for a in source]


def pack(motions):
    """-> (tables float64 [M,5,Lmax] zero padded, lengths int32 [M])"""
    lengths = np.asarray([m.shape[1] for m in motions], dtype=np.int32)
    out = np.zeros((len(motions), 5, int(lengths.max())), dtype=np.float64)
    for i, m in enumerate(motions):
        if m.shape[0] != 5:
            raise ValueError("motion %d: expected 5 rows" % i)
        out[i, :, :m.shape[1]] = m
    return out, lengths
