"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/*.npz by running the UNMODIFIED reference
(/root/reference, imported through oracle/ref_loader.py + oracle/pyg_shim) on CPU in fp32 and
fp64.  Run in the build container:   python oracle/make_golden.py
The fixtures travel to the GPU box; /root/reference does not.

Each fixture stores the reference module's state_dict (`sd/<key>`), the inputs, and the
reference's outputs (`*_f32` from the fp32 run, `*_f64` from module.double()).
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import ref_loader  # noqa: E402

ref_loader.load()
from sgnn.single_scale.learned_simulator import LearnedSimulator  # noqa: E402
from sgnn.single_scale import evaluate as ref_evaluate  # noqa: E402
from sgnn.multi_scale.multi_scale_simulator import MultiScaleSimulator  # noqa: E402
from sgnn.multi_scale.multi_scale_graph import MultiScaleGraph, MultiScaleConfig  # noqa: E402
from sgnn import noise_utils as ref_noise  # noqa: E402

from sgnn_b200 import synthetic  # noqa: E402  (input generator only; no product compute)

OUT = os.path.join(ROOT, "tests", "golden")


def _np(t):
    return t.detach().cpu().numpy()


def _stats_to(stats, dtype):
    return {k: {kk: vv.to(dtype) for kk, vv in v.items()} for k, v in stats.items()}


def _pack_sd(sd):
    return {"sd/" + k: _np(v) for k, v in sd.items()}


def _pack_stats(stats):
    return {f"stats/{k}/{kk}": _np(vv) for k, v in stats.items() for kk, vv in v.items()}


def _store_grads(out, tag, sim):
    """fp64 gradients are the truth (stored as fp32 to keep the fixtures small); of the fp32
    run only the per-parameter max |fp32 - fp64| gap is kept (tolerance calibration)."""
    for k, p in sim.named_parameters():
        if p.grad is None:
            continue
        if tag == "f32":
            out[f"_g32/{k}"] = p.grad.detach().double()
        else:
            out[f"grad_f64/{k}"] = _np(p.grad.float())
            g32 = out.pop(f"_g32/{k}")
            out[f"grad_gap/{k}"] = np.array([(g32 - p.grad).abs().max().item(),
                                             p.grad.abs().max().item()])


def single_scale_fixture(name, nx, ny, spacing, radius, S, H, L, nmlp, nsteps, batch_counts=None,
                         ntypes=1, emb=9, seed=1):
    traj = synthetic.taylor_trajectory(nx, ny, spacing, nframes=S + nsteps, seed=0)
    pos, types, strains = traj["positions"], traj["particle_types"], traj["strains"]
    if ntypes > 1:
        types = (torch.arange(pos.shape[0]) % ntypes).to(torch.int64)
    meta = synthetic.metadata_from(pos)
    noise_std = 0.02 * spacing / 0.5
    stats = synthetic.normalization_stats(meta, noise_std)
    nnode_in = (S - 1) * 2 + 1 + (emb if ntypes > 1 else 0)

    def make(dtype):
        torch.manual_seed(seed)
        sim = LearnedSimulator(2, nnode_in, 3, H, L, nmlp, H, radius, _stats_to(stats, dtype),
                               ntypes, emb, device="cpu")
        return sim.to(dtype)

    sim32 = make(torch.float32)
    sim64 = make(torch.float64)
    out = {"meta": np.array([nx, ny, S, H, L, nmlp, nsteps, ntypes, emb], dtype=np.int64),
           "spacing_radius": np.array([spacing, radius], dtype=np.float64),
           "positions": _np(pos), "particle_types": _np(types), "strains": _np(strains)}
    out.update(_pack_sd(sim32.state_dict()))
    out.update(_pack_stats(stats))

    # ---- rollout through the reference's own evaluate.rollout (evaluate.py:52) ----
    n = pos.shape[0]
    for tag, sim, dtype in (("f32", sim32, torch.float32), ("f64", sim64, torch.float64)):
        sim.eval()
        with torch.no_grad():
            # edge sets per step, recorded by re-running the connectivity on the window
            cur = pos[:, :S].to(dtype)
            eis, preds, sts = [], [], []
            for step in range(nsteps):
                nf, ei, ef = sim._encoder_preprocessor(cur, [torch.tensor(n)], types)
                nxt, st = sim.predict_positions(cur, [torch.tensor(n)], types)
                eis.append(_np(ei))
                preds.append(_np(nxt))
                sts.append(_np(st))
                cur = torch.cat([cur[:, 1:], nxt[:, None, :]], dim=1)
            res = ref_evaluate.rollout(sim, pos.to(dtype), types, torch.tensor(n),
                                       strains.to(dtype), nsteps, 2, "cpu", S, "autoregressive")
            assert np.array_equal(res["predicted_rollout"], np.stack(preds))
        out[f"rollout_{tag}"] = res["predicted_rollout"]
        out[f"rollout_strain_{tag}"] = res["predicted_strain"]
        if tag == "f32":
            for s, ei in enumerate(eis):
                out[f"edge_index/{s}"] = ei
            out["rmse_position_f32"] = res["rmse_position"]
            # first-step internals: features and network output
            nf, ei, ef = sim._encoder_preprocessor(pos[:, :S], [torch.tensor(n)], types)
            out["node_features_f32"], out["edge_features_f32"] = _np(nf), _np(ef)
            out["epd_out_f32"] = _np(sim._encode_process_decode(nf, ei, ef))

    # ---- one training step (train.py:230-280): noise, predict_accelerations, loss, grads ----
    counts = batch_counts or [n]
    if batch_counts:
        # ragged batch: cut examples of different sizes out of the lattice (rows of ny particles)
        parts, tps, nxts, nstr = [], [], [], []
        off = 0
        for c in counts:
            parts.append(pos[off:off + c, :S]); tps.append(types[off:off + c])
            nxts.append(pos[off:off + c, S]); nstr.append(strains[S, off:off + c])
            off += c
        bpos, btypes = torch.cat(parts), torch.cat(tps)
        bnext, bstrain = torch.cat(nxts), torch.cat(nstr)
    else:
        bpos, btypes, bnext, bstrain = pos[:, :S], types, pos[:, S], strains[S]
    torch.manual_seed(2)
    noise = ref_noise.get_random_walk_noise_for_position_sequence(bpos, noise_std)
    out["train/positions"], out["train/noise"] = _np(bpos), _np(noise)
    out["train/next_position"], out["train/next_strain"] = _np(bnext), _np(bstrain)
    out["train/counts"] = np.array(counts, dtype=np.int64)
    out["train/particle_types"] = _np(btypes)
    for tag, sim, dtype in (("f32", sim32, torch.float32), ("f64", sim64, torch.float64)):
        sim.train()
        sim.zero_grad()
        pa, ta, ps = sim.predict_accelerations(bnext.to(dtype), noise.to(dtype), bpos.to(dtype),
                                               torch.tensor(counts), btypes)
        loss = (((pa - ta) ** 2).sum(-1) + (ps - bstrain.to(dtype)) ** 2).mean()
        loss.backward()
        out[f"train/pred_acc_{tag}"], out[f"train/target_acc_{tag}"] = _np(pa), _np(ta)
        out[f"train/pred_strain_{tag}"] = _np(ps)
        out[f"train/loss_{tag}"] = _np(loss)
        _store_grads(out, tag, sim)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print(name, "N", n, "E0", out["edge_index/0"].shape[1], "loss", float(out["train/loss_f32"]))


def multi_scale_fixture(name, positions0, S, nnode_extra, H, nedge_out, L, nmlp, num_scales,
                        window, mult, ntypes, emb, nsteps, seed=1, jitter=0.01):
    """positions0 [N,2]: frame-0 lattice (exact coordinates, needed by the hierarchy sampling)."""
    n = positions0.shape[0]
    g = torch.Generator().manual_seed(0)
    frames = [positions0.to(torch.float64)]
    for t in range(1, S + nsteps):
        frames.append(frames[0] + torch.tensor([-0.05 * t, 0.0], dtype=torch.float64)
                      + jitter * torch.randn(n, 2, generator=g, dtype=torch.float64))
    pos = torch.stack(frames, dim=1).to(torch.float32).contiguous()
    types = (torch.arange(n) % ntypes).to(torch.int64)
    meta = synthetic.metadata_from(pos)
    stats = synthetic.normalization_stats(meta, 0.02)
    nnode_in = (S - 1) * 2 + 1 + (emb if ntypes > 1 else 0)

    def make(dtype):
        torch.manual_seed(seed)
        sim = MultiScaleSimulator(2, nnode_in, 3, nedge_out, H, L, nmlp, _stats_to(stats, dtype),
                                  ntypes, emb, num_scales=num_scales, window_size=window,
                                  radius_multiplier=mult, device="cpu")
        return sim.to(dtype)

    graph = MultiScaleGraph(MultiScaleConfig(num_scales, window, mult)).create_all_edges(pos[:, 0])
    out = {"meta": np.array([S, H, nedge_out, L, nmlp, num_scales, window, ntypes, emb, nsteps],
                            dtype=np.int64),
           "mult": np.array([mult], dtype=np.float64),
           "positions": _np(pos), "particle_types": _np(types),
           "g2m": _np(graph["grid2mesh_edges"]), "m2m": _np(graph["mesh2mesh_edges"]),
           "m2g": _np(graph["mesh2grid_edges"])}
    for s, h in graph["graph_hierarchy"].items():
        out[f"hier/{s}/sampling_indices"] = _np(h["sampling_indices"])
        out[f"hier/{s}/spacing"] = np.array([h["spacing"]])
    out.update(_pack_stats(stats))
    for tag, dtype in (("f32", torch.float32), ("f64", torch.float64)):
        sim = make(dtype)
        if tag == "f32":
            out.update(_pack_sd(sim.state_dict()))
        sim.set_static_graph(graph)
        sim.eval()
        with torch.no_grad():
            cur = pos[:, :S].to(dtype)
            preds, sts = [], []
            for _ in range(nsteps):
                nxt, st = sim.predict_positions(cur, torch.tensor([n]), types)
                preds.append(_np(nxt)); sts.append(_np(st))
                cur = torch.cat([cur[:, 1:], nxt[:, None, :]], dim=1)
            if tag == "f32":
                nf, eidx, ef = sim._encoder_preprocessor(pos[:, :S], torch.tensor([n]), types)
                out["node_features_f32"] = _np(nf)
                for k in ("g2m", "m2m", "m2g"):
                    out[f"edge_features_f32/{k}"] = _np(ef[k])
        out[f"rollout_{tag}"] = np.stack(preds)
        out[f"rollout_strain_{tag}"] = np.stack(sts)
        # training step
        torch.manual_seed(2)
        noise = ref_noise.get_random_walk_noise_for_position_sequence(pos[:, :S], 0.02)
        sim.train(); sim.zero_grad()
        pa, ta, ps = sim.predict_accelerations(pos[:, S].to(dtype), noise.to(dtype),
                                               pos[:, :S].to(dtype), torch.tensor([n]), types)
        strain_t = torch.zeros(n, dtype=dtype)
        loss = (((pa - ta) ** 2).sum(-1) + (ps - strain_t) ** 2).mean()
        loss.backward()
        out["train/noise"] = _np(noise)
        out[f"train/pred_acc_{tag}"], out[f"train/target_acc_{tag}"] = _np(pa), _np(ta)
        out[f"train/pred_strain_{tag}"], out[f"train/loss_{tag}"] = _np(ps), _np(loss)
        _store_grads(out, tag, sim)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print(name, "N", n, "edges", out["g2m"].shape[1], out["m2m"].shape[1], out["m2g"].shape[1])


def radius_fixture():
    """Edge sets obtained through the reference's own call site
    (LearnedSimulator._compute_graph_connectivity, learned_simulator.py:70-124)."""
    stats = {"velocity": {"mean": torch.zeros(2), "std": torch.ones(2)},
             "acceleration": {"mean": torch.zeros(2), "std": torch.ones(2)}}
    sim = LearnedSimulator(2, 21, 3, 8, 1, 1, 8, 1.0, stats, 1, 9)
    out = {}
    rng = np.random.default_rng(0)
    cases = {
        # exactly representable lattice: d^2 == r^2 pairs must be EXCLUDED (strict <)
        "lattice_exact": (synthetic.taylor_bar_lattice(16, 10, 0.5), [160], 1.5),
        "lattice_trunc": (synthetic.taylor_bar_lattice(20, 12, 0.5), [240], 15.0),
        "random": (torch.from_numpy(rng.uniform(0, 20, (500, 2)).astype(np.float32)), [500], 1.7),
        "ragged_batch": (torch.from_numpy(rng.uniform(0, 8, (403, 2)).astype(np.float32)),
                         [1, 200, 2, 200], 1.1),
        "dim3": (torch.from_numpy(rng.uniform(0, 6, (400, 3)).astype(np.float32)), [400], 1.2),
        "single": (torch.zeros(1, 2), [1], 0.6),
        "duplicates": (torch.from_numpy(np.repeat(rng.uniform(0, 2, (30, 2)), 3, 0)
                                        .astype(np.float32)), [90], 0.7),
    }
    for name, (pos, counts, r) in cases.items():
        a, b = sim._compute_graph_connectivity(pos, torch.tensor(counts), r)
        out[name + "/pos"] = _np(pos)
        out[name + "/counts"] = np.array(counts, dtype=np.int64)
        out[name + "/r"] = np.array([r])
        out[name + "/edge_index"] = np.stack([_np(a), _np(b)])
        print("radius", name, out[name + "/edge_index"].shape)
    np.savez_compressed(os.path.join(OUT, "radius_cases.npz"), **out)


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(1)  # run-to-run identical reduction order
    radius_fixture()
    # BASELINE model (H=64, L=5, S=11, nmlp=1) on a small bar, truncation-binding spacing B
    single_scale_fixture("single_h64", nx=20, ny=12, spacing=5.0, radius=15.0, S=11, H=64, L=5,
                         nmlp=1, nsteps=5)
    # ragged batch, 2 particle types (embedding path), nmlp_layers=2, sparse regime A
    single_scale_fixture("single_h32_batch", nx=16, ny=10, spacing=12.5, radius=15.0, S=6, H=32,
                         L=2, nmlp=2, nsteps=3, batch_counts=[50, 70, 40], ntypes=2, emb=9)
    # the reference's own simulator test fixture (test_multi_scale_simulator.py:27-96)
    x = torch.linspace(-2, 2, 5)
    xx, yy = torch.meshgrid(x, x, indexing="ij")
    multi_scale_fixture("multi_ref_fixture", torch.stack([xx.flatten(), yy.flatten()], 1), S=6,
                        nnode_extra=16, H=64, nedge_out=32, L=3, nmlp=2, num_scales=3, window=3,
                        mult=2.0, ntypes=2, emb=16, nsteps=2)
    # config-4-shaped (H=128, nmlp=2, scales=2, window=2) on a small 0.5-spaced bar, L=1
    multi_scale_fixture("multi_h128", synthetic.taylor_bar_lattice(24, 16, 0.5), S=11,
                        nnode_extra=0, H=128, nedge_out=128, L=1, nmlp=2, num_scales=2, window=2,
                        mult=2.0, ntypes=1, emb=9, nsteps=2)
