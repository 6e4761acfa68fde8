"""Generates the committed golden vectors of tests/golden/ from the CPU oracle.

    python tests/golden/make_golden.py

The reference ships no golden vectors. cfg1_ekf.npz, cfg2_ukf.npz and cfg3_ekf_sample.npz are OUTPUTS OF THE
REFERENCE ITSELF run here: oracle/_ref (ekf.cpp / ukf.cpp compiled unmodified against the Eigen / ROS stand-ins of
oracle/ref_shim; cfg3 with MAX_LANDMARK_COUNT lifted to 64 at build time) -- this script runs the compiled reference
and the restated oracle side by side, asserts they are bit-equal, and stores X, P, N from the reference plus the
association indices and LLT failure columns only the oracle can report. large_ekf_l6.npz (the row-subset large-map
step, a formulation the reference does not have) stays an oracle-generated regression pin. Each .npz holds the
complete replay inputs, so neither the tests nor the GPU box depend on the replay generator or on /root/reference.

  cfg1_ekf.npz  single EKF-SLAM filter, landmarks.yaml map, 1000 messages, seed 1 (BASELINE config 1)
  cfg2_ukf.npz  single UKF-SLAM filter, same replay                                 (BASELINE config 2)
     inputs : x, y, vx, az (f64 [T]); qw, qz (f64 [T]: orientation (w, 0, 0, z)); yaw = quat2euler of it, dt
              (f32 [T]); range, bearing (f32 [T][8])
     outputs: steps (the checked message indices), N[k], X[k][29], P[k][29][29] (zero padded),
              assoc[k][8], llt_fail[T] (UKF: first failing pivot per message, -1 = factorised)
  cfg3_ekf_sample.npz  filters 0..3 of the cfg3 batch (seeds 0x5EED0000 + id), 40 messages: final X, P, N
  large_ekf_l6.npz     large-map EKF row-subset step, 6x6 landmarks (n = 75), one message: X, P after
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

import awesomeslam_b200 as A  # noqa: E402  (replay generator: host code of the product library)
from oracle import oracle as O  # noqa: E402
from oracle import ref as R  # noqa: E402

STEPS = [0, 9, 10, 11, 12, 100, 500, 999]
LD = 29


def single(kind, path):
    lmx, lmy = A.map_cfg1()
    T = 1000
    rp = A.Replay(lmx, lmy, 1, T, seed0=A.replay.SEED_CFG1)
    o = O.Filter(kind)
    node = R.Node(kind)
    Ns, Xs, Ps, As = [], [], [], []
    fails = np.full(T, -1, np.int32)
    for t in range(T):
        o.set_observations(rp.range[t, 0], rp.bearing[t, 0])
        o.step(rp.x[t, 0], rp.y[t, 0], rp.yaw[t, 0], rp.vx[t, 0], rp.az[t, 0], rp.dt[t, 0])
        node.set_time(float(t + 1))
        node.landmarks(rp.range[t, 0].astype(np.float64), rp.bearing[t, 0].astype(np.float64))
        node.odom(rp.x[t, 0], rp.y[t, 0], rp.qw[t, 0], 0.0, 0.0, rp.qz[t, 0], rp.vx[t, 0], rp.az[t, 0])
        if kind == "ukf":
            fails[t] = o.llt_fail()
        if t in STEPS:
            X, Z, P = node.state()
            n = node.N
            Xo, Zo, Po = o.state()
            assert n == o.N and np.array_equal(X, Xo) and np.array_equal(P, Po) and np.array_equal(Z, Zo), t
            Xp = np.zeros(LD)
            Pp = np.zeros((LD, LD))
            Xp[:n] = X
            Pp[:n, :n] = P
            a = np.full(8, -2, np.int32)
            a[:] = o.assoc()
            Ns.append(n)
            Xs.append(Xp)
            Ps.append(Pp)
            As.append(a)
    np.savez_compressed(path, x=rp.x[:, 0], y=rp.y[:, 0], yaw=rp.yaw[:, 0], vx=rp.vx[:, 0], az=rp.az[:, 0],
                        dt=rp.dt[:, 0], qw=rp.qw[:, 0], qz=rp.qz[:, 0], range=rp.range[:, 0], bearing=rp.bearing[:, 0],
                        steps=np.array(STEPS),
                        N=np.array(Ns), X=np.array(Xs), P=np.array(Ps), assoc=np.array(As), llt_fail=fails)
    print(path, "final N", Ns[-1], "LLT failures", int(np.sum(fails >= 0)))


def cfg3_sample(path):
    lmx, lmy = A.map_cfg3()
    B, T, L = 4, 40, 16
    rp = A.Replay(lmx, lmy, B, T, seed0=A.replay.SEED_CFG3)
    cfg = O.default_config(max_state_dim=36)
    sec, X, P, N = O.run_batch("ekf", cfg, B, T, L, rp.x, rp.y, rp.yaw, rp.vx, rp.az, rp.dt, rp.range, rp.bearing,
                               1, 36)
    _, Xr, Pr, Nr = R.run_batch("ekf", B, T, L, rp.x, rp.y, rp.qw, rp.qz, rp.vx, rp.az, rp.range, rp.bearing, 1, 36,
                                cap64=True)
    assert np.array_equal(N, Nr) and np.array_equal(X, Xr) and np.array_equal(P, Pr)
    np.savez_compressed(path, x=rp.x, y=rp.y, yaw=rp.yaw, vx=rp.vx, az=rp.az, dt=rp.dt, qw=rp.qw, qz=rp.qz,
                        range=rp.range, bearing=rp.bearing, X=Xr, P=Pr, N=Nr)
    print(path, "N", N)


def large_l6(path):
    tx, ty, X0, G, d = A.large_map_state(6)
    P0 = A.dense_P(G, d)
    rng, brg, idx = A.large_observations(tx, ty, (0.0, 0.0, 0.0), m=8)
    X, P = X0.copy(), P0.copy()
    odom = (0.01, -0.01, 0.002, 0.05, 0.1, 1.0)
    q, assoc, a00, a10 = O.ekf_large_step(O.default_config(), X, P, 1.0, 0.0, *odom, rng, brg)
    np.savez_compressed(path, X0=X0, P0=P0, range=rng, bearing=brg, odom=np.array(odom), X=X, P_colmajor=P,
                        assoc=assoc, q=q, a=np.array([a00, a10]))
    print(path, "q", q, "assoc", assoc)


if __name__ == "__main__":
    O.build()
    R.build()
    single("ekf", os.path.join(HERE, "cfg1_ekf.npz"))
    single("ukf", os.path.join(HERE, "cfg2_ukf.npz"))
    cfg3_sample(os.path.join(HERE, "cfg3_ekf_sample.npz"))
    large_l6(os.path.join(HERE, "large_ekf_l6.npz"))
