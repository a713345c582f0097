"""CPU: host-side logic of the product — URDF/motion parsing, the C-ABI surface, sharding. No compute calls."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from conftest import GOLDEN, REPO

PKG = "a-deep-reinforcement-learning-enabled-soft-exoskeleton-for-parkinson-s-patients_b200"


def test_alias_and_package_import(exo):
    import importlib
    assert importlib.import_module(PKG).BatchedExoskeletonEnv is exo.BatchedExoskeletonEnv
    assert exo.ExoskeletonEnv_train.__init__.__code__.co_varnames[1:11] == (
        "reference_motion_file_num", "tremor_sequence", "tremor_amplitude_range", "first_harmonics_interval",
        "second_harmonics_interval", "max_force_shoulder", "max_force_elbow", "dr_actuator_end_pos_shift",
        "dr_actuator_range", "matrix_noise_fraction")          # Exoskeleton_env.py:38-48, same names and order


def test_geometry_matches_independent_urdf_parse(exo):
    """The product's parser (urdf.py) and the oracle's (minibullet.py) read the reference URDF independently."""
    tree = exo.urdf.load_geometry()
    d = np.load(os.path.join(GOLDEN, "ref_data_and_units.npz"))
    links = tree["links"]
    assert [l["name"] for l in links] == list(d["geo/names"])
    assert [l["parent"] for l in links] == list(d["geo/parent"])
    assert [l["joint_type"] for l in links] == list(d["geo/joint_type"])
    for key in ("origin_xyz", "origin_R", "axis", "com_xyz", "lower", "upper"):
        assert np.allclose(np.array([l[key] for l in links]), d["geo/" + key], atol=1e-15), key
    assert np.allclose(tree["base_pos"], d["geo/base_pos"])
    # link handles hard-coded by the reference (Exoskeleton_sim_pybullet.py:48-63) point at the right links
    names = [l["name"] for l in links]
    want = ["k11", "k12", "k21", "k22", "k31", "k32", "k41", "k42", "k51", "k52", "k61", "k62", "k71", "k72"]
    assert [names[i] for i in tree["act_link"]] == want
    assert names[tree["shoulder_link"]] == "auxlink1" and names[tree["elbow_link"]] == "auxlink3"


def test_urdf_parser_on_a_file(tmp_path, exo):
    urdf = tmp_path / "two.urdf"
    urdf.write_text("""<?xml version="1.0"?><robot name="r"><link name="b"/><link name="a"><inertial><origin xyz="0 0.5 0"/></inertial></link>
    <link name="c"/><joint type="revolute" name="j0"><axis xyz="0 0 1"/><limit lower="-1" upper="2"/><parent link="b"/><child link="a"/>
    <origin xyz="1 2 3" rpy="0 0 1.5707963267948966"/></joint>
    <joint type="prismatic" name="j1"><axis xyz="0 0 1"/><parent link="a"/><child link="c"/><origin xyz="0 0 1" rpy="0 0 0"/></joint></robot>""")
    t = exo.urdf.parse_urdf(str(urdf), base_position=(0, 0, 0))
    assert [l["name"] for l in t["links"]] == ["a", "c"] and [l["parent"] for l in t["links"]] == [-1, 0]
    assert np.allclose(np.array(t["links"][0]["origin_R"]).reshape(3, 3), [[0, -1, 0], [1, 0, 0], [0, 0, 1]], atol=1e-15)
    assert t["links"][0]["com_xyz"] == [0.0, 0.5, 0.0] and t["links"][0]["lower"] == -1 and t["links"][1]["joint_type"] == 1


def test_motion_tables(tmp_path, exo):
    d = np.load(os.path.join(GOLDEN, "ref_data_and_units.npz"))
    ms = exo.motions.load_motions()
    assert [m.shape[1] for m in ms] == [232, 316, 282, 295, 270, 347, 266, 273]
    for i, m in enumerate(ms):
        assert np.array_equal(m, d["motion/%d" % i])          # bit-identical to the reference's read_env_texts
    # text parser: 17 columns, first five kept; ragged / short rows rejected
    f = tmp_path / "ref_motion_0.txt"
    rows = np.arange(3 * 17, dtype=np.float64).reshape(3, 17) / 7
    f.write_text("\n".join(" ".join(repr(float(x)) for x in r) for r in rows) + "\n\n")
    got = exo.motions.read_motion_file(str(f))
    assert got.shape == (5, 3) and np.array_equal(got, rows[:, :5].T)
    f.write_text("1 2 3\n")
    with pytest.raises(ValueError):
        exo.motions.read_motion_file(str(f))
    tables, lengths = exo.motions.pack(ms)
    assert tables.shape == (8, 5, 347) and list(lengths) == [m.shape[1] for m in ms] and tables[0, :, 232:].sum() == 0


def _header_functions():
    text = open(os.path.join(REPO, "include", "exo_env.h")).read() + open(os.path.join(REPO, "include", "exo_lap.h")).read() + \
        open(os.path.join(REPO, "include", "exo_policy.h")).read() + open(os.path.join(REPO, "include", "exo_td7.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(exo_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(exo):
    lib = exo.load_library()                                  # dlopen works without a GPU
    declared = _header_functions()
    assert len(declared) >= 16
    for name in declared:
        assert hasattr(lib, name), name
    assert sorted(exo._capi.SIGNATURES) == declared           # the ctypes table binds exactly the header's surface
    assert lib.exo_abi_version() == 1


def test_ctypes_structs_match_the_header(tmp_path, exo):
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "exo_env.h"\nint main(){printf("%zu %zu %zu %d %zu %zu\\n",'
                   'sizeof(ExoConfig),sizeof(ExoGeometry),sizeof(ExoInfo),(int)EXO_N_FIELDS,offsetof(ExoConfig,link_state_lag),'
                   'offsetof(ExoGeometry,act_link));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-I", os.path.join(REPO, "include"), str(src), "-o", str(exe)])
    sizes = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    c = exo._capi
    assert sizes == [ctypes.sizeof(c.ExoConfig), ctypes.sizeof(c.ExoGeometry), ctypes.sizeof(c.ExoInfo),
                     len(c.FIELDS), c.ExoConfig.link_state_lag.offset, c.ExoGeometry.act_link.offset]


def test_no_gpu_means_loud_failure(exo):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(exo.ExoError):
        exo.BatchedExoskeletonEnv([0], [[1, 1, 1, 1, 0, 0, 0]], [0], tremor_amplitude_range=[1, 1], first_harmonics_interval=[4, 6],
                                  second_harmonics_interval=[8, 10], max_force_shoulder=40, max_force_elbow=20)
    # and the C-ABI itself refuses rather than computing on the host
    lib = exo.load_library()
    h = ctypes.c_void_p()
    cfg = exo._capi.ExoConfig()
    geo = exo.urdf.to_struct(exo.urdf.load_geometry())
    rc = lib.exo_create(ctypes.byref(cfg), ctypes.byref(geo), 4, 0, ctypes.byref(h))
    assert rc < 0 and lib.exo_last_error()


def test_product_never_imports_the_oracle():
    """The oracle is a checker: nothing under the package (or its C sources) may reference it."""
    bad = []
    for root, _, files in os.walk(os.path.join(REPO, PKG)):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(root, f), errors="ignore").read()
                if re.search(r"^\s*(import|from)\s+(oracle|minibullet)|exo_oracle|libexo_oracle|#include\s+\"[^\"]*oracle", text, flags=re.M):
                    bad.append(f)
    assert not bad, bad


def test_shard_ranges_partition_the_population(exo):
    sh = exo.sharding
    for n, w in ((1_048_576, 8), (65_536, 1), (1000, 3), (7, 8)):
        spans = [sh.shard_range(n, r, w) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n and all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
    m, s, seeds = sh.benchmark_population(65_536, 0, 1)
    assert np.array_equal(np.bincount(m), np.full(8, 8192))                                   # cfg 3: 8192 envs per motion
    assert all(len(set(m[b:b + 1024])) == 1 for b in range(0, 65_536, 1024))                   # in contiguous 1024-env blocks
    assert len({tuple(x) for x in s}) == 15 and s[:, :4].sum(1).min() >= 1 and s[:, 4:].sum() == 0
    # global ids/seeds do not depend on the number of ranks
    parts = [sh.benchmark_population(4096, r, 4) for r in range(4)]
    assert np.array_equal(np.concatenate([p[2] for p in parts]), np.arange(4096))
    assert np.array_equal(np.concatenate([p[1] for p in parts]), sh.benchmark_population(4096, 0, 1)[1])
    # ... nor does anything else about global env g: motion, tremor type and seed of the concatenated shards are those of
    # the one-rank population for every world size (BASELINE config 4: 1M envs over 2/4/8 GPUs)
    for n in (1_048_576, 65_536, 5000):
        one = sh.benchmark_population(n, 0, 1)
        for w in (2, 4, 8):
            parts = [sh.benchmark_population(n, r, w) for r in range(w)]
            for k in range(3):
                assert np.array_equal(np.concatenate([p[k] for p in parts]), one[k]), (n, w, k)


def test_reference_stream_reproduces_the_shared_numpy_stream_bit_for_bit(exo):
    """8 reference envs built after ONE np.random.seed, reset in order, warm-up actions from the same stream
    (Exoskeleton_agent_train.py:51-71,110-131): exo_b200.reference_stream restates the draw ORDER on the host. Every
    RNG-derived quantity of the golden (tests/golden/ref_shared_stream.npz, reference Python) must come out bit-identical:
    tremor tables, I/D/S, shifts, force maxima, magnitudes and the actions themselves."""
    from conftest import GOLDEN, case_settings
    rsm = exo.reference_stream
    g = np.load(os.path.join(GOLDEN, "ref_shared_stream.npz"))
    st = case_settings(g, "S0")
    L = [int(g["S%d/meta_max_count" % i]) for i in range(8)]
    seqs = [g["S%d/meta_seq" % i] for i in range(8)]
    stream = rsm.ReferenceStream(int(g["S0/meta_seed"]))
    for i in range(8):
        stream.constructor(L[i], seqs[i], **st)                 # the constructors' own initialize_movement draws
    eps = [stream.initialize_movement(L[i], seqs[i], **st) for i in range(8)]        # the first reset() of each env, in order
    for i in range(8):
        c = "S%d" % i
        assert np.array_equal(eps[i]["tremor_table"], g[c + "/ep_tt"][0])
        assert eps[i]["tremor_mag"] == float(g[c + "/ep_mag"][0])
        for k, key in (("I", "ep_I"), ("D", "ep_D"), ("S", "ep_S"), ("shift", "ep_shift")):
            assert np.array_equal(eps[i][k], g[c + "/" + key][0]), (i, k)
        assert np.array_equal(eps[i]["force_max"], g[c + "/ep_fmax"][0])
    done = [False] * 8
    n = [0] * 8
    for t in range(int(g["S/n_steps"])):
        for i in range(8):
            if not done[i]:
                a = stream.warmup_action().astype(np.float32).astype(np.float64)
                assert np.array_equal(a, g["S%d/action" % i][n[i]])
        for i in range(8):
            if not done[i]:
                done[i] = bool(g["S%d/done" % i][n[i]])
                n[i] += 1
    # the exact propagator of the host matches the kernel's arithmetic contract: G T = scipy RK45 solution to its tolerance
    d = np.load(os.path.join(GOLDEN, "ref_data_and_units.npz"))
    G = rsm.propagator(d["mat/I"], d["mat/D"], d["mat/S"])
    for T, q in zip(d["ode/T"], d["ode/q"]):
        assert np.max(np.abs(G @ T - q)) <= 1e-3 * max(np.max(np.abs(q)), 1e-9)


def test_episode_suppression_log_matches_the_training_scripts_numpy_statistics(exo):
    """Per-env medians / standard deviations / means of the training script (Simulation/Exoskeleton_agent_train.py:150-205
    per step, :239-265 per episode; VERDICT r01 'missing' 7): exo_b200.metrics.EpisodeSuppressionLog against a literal numpy
    restatement of the script's loop on the same synthetic info."""
    import torch
    rs = np.random.RandomState(4)
    n, T = 5, 37
    lens = [37, 30, 37, 12, 25]
    log = exo.metrics.EpisodeSuppressionLog(n, T, "cpu", disregard=True)
    red_full, ampl_full, tot_full = [], [], []
    when, when_ampl = np.zeros((n, 2)), np.zeros((n, 2))
    for t in range(T):
        info = {k: rs.uniform(-2, 2, (n, 7)) for k in ("torque_val", "tremor_torque_val", "ampl_val", "tremor_ampl_val")}
        info["tremor_torque_val"][:, 4:] = 0.0                        # axes without tremor: 0/0 -> nan -> 0
        info["torque_val"][:, 4:] = 0.0
        info["amp_metric"] = rs.uniform(-50, 50, n)
        live = np.array([t < l - 3 for l in lens])
        log.add({k: torch.from_numpy(v) for k, v in info.items()}, torch.from_numpy(live))
        red_ep, ampl_ep, tot_ep = np.zeros((n, 7)), np.zeros((n, 7)), np.zeros(n)
        for i in range(n):                                            # the script, line by line
            if not live[i]:
                continue
            with np.errstate(all="ignore"):
                tr = (abs(info["torque_val"][i]) - abs(info["tremor_torque_val"][i])) / abs(info["tremor_torque_val"][i]) * 100
                ta = (abs(info["ampl_val"][i]) - abs(info["tremor_ampl_val"][i])) / abs(info["tremor_ampl_val"][i]) * 100
            tr = np.nan_to_num(tr, nan=0, posinf=0, neginf=0)
            ta = np.nan_to_num(ta, nan=0, posinf=0, neginf=0)
            tot = info["amp_metric"][i]
            when[i, 0] += np.sum(tr[:4] >= 0); when[i, 1] += np.sum(tr[:4] < 0)
            if tot < 0:
                when_ampl[i, 1] += 1; tot_ep[i] = tot
            else:
                when_ampl[i, 0] += 1
            tr[tr > 0] = 0; ta[ta > 0] = 0
            if len(np.nonzero(tr)[0]) > 0:
                red_ep[i] = tr
            if len(np.nonzero(ta)[0]) > 0:
                ampl_ep[i] = ta
        red_full.append(red_ep); ampl_full.append(ampl_ep); tot_full.append(tot_ep)
    red_np, ampl_np, tot_np = np.array(red_full), np.array(ampl_full), np.array(tot_full)
    s = {k: v.numpy() for k, v in log.summary(lens).items()}
    for i in range(n):
        a, b = red_np[:lens[i], i], ampl_np[:lens[i], i]
        assert np.allclose(s["tremor_avg"][i], a.mean(0), atol=1e-12) and np.allclose(s["tremor_median"][i], np.median(a, 0), atol=1e-12)
        assert np.allclose(s["tremor_std"][i], a.std(0), atol=1e-12)
        assert np.allclose(s["tremor_ampl_avg"][i], b.mean(0), atol=1e-12) and np.allclose(s["tremor_ampl_median"][i], np.median(b, 0), atol=1e-12)
        assert np.allclose(s["tremor_ampl_std"][i], b.std(0), atol=1e-12)
        assert abs(s["tremor_sup_avg"][i] - a[a != 0].mean()) < 1e-12 and abs(s["tremor_sup_median"][i] - np.median(a[a != 0])) < 1e-12
        assert abs(s["tremor_ampl_sup_avg"][i] - b[b != 0].mean()) < 1e-12
        assert abs(s["tremor_ampl_total_avg"][i] - tot_np[:lens[i], i].mean()) < 1e-12
        assert abs(s["tremor_occurrence_pct"][i] - when[i, 1] / when[i].sum() * 100) < 1e-12
        assert abs(s["tremor_ampl_occurrence_pct"][i] - when_ampl[i, 1] / when_ampl[i].sum() * 100) < 1e-12
