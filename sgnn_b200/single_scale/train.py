"""Entry points of the single-scale simulator: the callers on either side of the hot path
(reference sgnn/single_scale/train.py: load_config :29, predict :53, load_model :169, train :185,
_get_simulator :431, main :494).  Same YAML keys (sgnn/single_scale/config.yaml), same flags
(--config --mode --model_file --log), same checkpoint names (model-best-{step:06}.pt,
train_state-best-{step:06}.pt with {'optimizer_state', 'global_train_state': {'step',
'lowest_eval_loss'}}) and the same rollout pickle.  Differences, all on the host side:

  * the data path is device-resident (sgnn_b200/data.py): no per-step host work or H2D copy;
  * training is data-parallel when launched under torchrun (one process per GPU, flat-gradient
    NCCL allreduce, sgnn_b200/training.py); rank 0 validates and writes checkpoints;
  * a resumed run continues from the restored step and learning rate (the reference resets `step`
    to 0 after loading, train.py:221 -- SURVEY.md 8(f) rank 4);
  * the optimisation step runs as ONE CUDA graph of hand-written kernels (sgnn_b200/train_engine.py: weight
    images, neighbour search, forward, loss, backward, gradient allreduce, Adam) whenever the model is one the
    engine covers and every batch has the same particle count; otherwise through torch.autograd (train_ops.py);
  * the loss scalar is read back every `log_every` steps only, not every step;
  * Weights & Biases logging is not wired (no network); `log` is accepted and ignored.
"""
import argparse
import json
import os
import pickle
import time
from pathlib import Path
from typing import Optional

import torch
import torch.distributed as dist
import yaml

from .. import data as data_mod
from .. import train_engine, training
from . import evaluate, learned_simulator


def load_config(config_path):
    """YAML -> dict, with the reference's defaults for keys it post-processes (train.py:29-50)."""
    with open(config_path) as f:
        config = yaml.safe_load(f)
    if config.get("model_file") in ("None", "none", ""):
        config["model_file"] = None
    config.setdefault("train_state_file", None)
    config.setdefault("inference_mode", "autoregressive")
    config.setdefault("loss_weight_position", 1.0)
    config.setdefault("loss_weight_strain", 1.0)
    config.setdefault("particle_type_embedding_size", 16)
    config.setdefault("run_name", "run")
    return config


def _get_simulator(metadata: dict, acc_noise_std: float, vel_noise_std: float, device, config: dict
                   ) -> learned_simulator.LearnedSimulator:
    """train.py:431-491: std = sqrt(std^2 + noise_std^2); nnode_in = (S-1)*dim + 1 (+ embedding when
    there are several particle types); nedge_in = dim + 1; nmlp_layers = 1."""
    stats = {
        "acceleration": {"mean": torch.FloatTensor(metadata["acc_mean"]).to(device),
                         "std": torch.sqrt(torch.FloatTensor(metadata["acc_std"]) ** 2 + acc_noise_std ** 2).to(device)},
        "velocity": {"mean": torch.FloatTensor(metadata["vel_mean"]).to(device),
                     "std": torch.sqrt(torch.FloatTensor(metadata["vel_std"]) ** 2 + vel_noise_std ** 2).to(device)},
    }
    ntypes = metadata.get("num_particle_types", 1)
    nnode_in = (config["input_sequence_length"] - 1) * config["dim"] + 1
    if ntypes > 1:
        nnode_in += config["particle_type_embedding_size"]
    return learned_simulator.LearnedSimulator(
        particle_dimensions=config["dim"], nnode_in=nnode_in, nedge_in=config["dim"] + 1,
        latent_dim=config["hidden_dim"], nmessage_passing_steps=config["layers"], nmlp_layers=1,
        mlp_hidden_dim=config["hidden_dim"], connectivity_radius=config["connection_radius"],
        normalization_stats=stats, nparticle_types=ntypes,
        particle_type_embedding_size=config["particle_type_embedding_size"], device=device)


def _rollout_one(simulator, traj, metadata, config, device):
    nsteps = metadata["sequence_length"] - config["input_sequence_length"]
    out = evaluate.rollout(simulator, traj["positions"].to(device), traj["particle_type"].to(device),
                           traj["n_particles_per_example"].to(device), traj["strains"].to(device), nsteps,
                           config["dim"], device, config["input_sequence_length"], config["inference_mode"])
    out["metadata"] = metadata
    return out


def validate(simulator, metadata: dict, config: dict, device, split: str = "valid"):
    """Mean over trajectories of final rmse_position + rmse_strain (train.py:303-357)."""
    was_training = simulator.training
    simulator.eval()
    total, outs = [], []
    with torch.no_grad():
        for traj in data_mod.get_data_loader_by_trajectories(str(Path(config["data_path"]) / f"{split}.npz"), device):
            out = _rollout_one(simulator, traj, metadata, config, device)
            total.append(float(out["rmse_position"][-1] + out["rmse_strain"][-1]))
            outs.append(out)
    if was_training:
        simulator.train()
    return sum(total) / len(total), outs


def predict(simulator, metadata: dict, device, config: dict):
    """Rollouts over the valid (mode 'valid') or test (mode 'rollout') split; mode 'rollout' writes
    one pickle per case under output_path/run_name (train.py:53-166)."""
    model_path = Path(config["model_path"]) / config["run_name"] / config["model_file"]
    simulator.load(str(model_path))
    simulator.to(device)
    simulator.eval()
    split = "test" if config["mode"] == "rollout" else "valid"
    mean_loss, outs = validate(simulator, metadata, config, device, split)
    if config["mode"] == "rollout":
        save_dir = Path(config["output_path"]) / config["run_name"]
        save_dir.mkdir(parents=True, exist_ok=True)
        names = metadata.get("file_test") or [f"rollout_{i}.npz" for i in range(len(outs))]
        for i, out in enumerate(outs):
            out["case_name"] = names[i].replace(".npz", "")
            with open(save_dir / f"{out['case_name']}.pkl", "wb") as f:
                pickle.dump(out, f)
    print(f"Mean loss on rollout prediction: {mean_loss}")
    return mean_loss


def load_model(simulator, device, config: dict):
    """(simulator, step, optimizer, lowest_eval_loss) from model_file + train_state_file under
    model_path/run_name (utils/checkpoint_utils.py:13-44)."""
    d = Path(config["model_path"]) / config["run_name"]
    mp, sp = d / config["model_file"], d / str(config["train_state_file"])
    if not (mp.exists() and sp.exists()):
        raise FileNotFoundError(f"Specified model_file {mp} and train_state_file {sp} not found.")
    simulator.load(str(mp))
    simulator.to(device)                       # before the optimizer: load_state_dict casts its state to the parameters' device
    state = torch.load(str(sp), map_location="cpu")
    optimizer = training.make_optimizer(simulator.parameters(), 1e-3, device)
    optimizer.load_state_dict(state["optimizer_state"])
    gts = state["global_train_state"]
    return simulator, int(gts["step"]), optimizer, float(gts.get("lowest_eval_loss", float("inf")))


def save_checkpoint(simulator, optimizer, step: int, lowest_eval_loss: float, save_dir: Path):
    save_dir.mkdir(parents=True, exist_ok=True)
    simulator.save(str(save_dir / f"model-best-{step:06}.pt"))
    torch.save({"optimizer_state": optimizer.state_dict(),
                "global_train_state": {"step": step, "lowest_eval_loss": lowest_eval_loss}},
               str(save_dir / f"train_state-best-{step:06}.pt"))


def train(simulator, metadata: dict, device, config: dict, log_every: int = 10, seed: int = 0):
    """Training loop (train.py:185-428).  Returns (last step, lowest validation loss)."""
    rank = dist.get_rank() if dist.is_initialized() else 0
    world = dist.get_world_size() if dist.is_initialized() else 1
    simulator.to(device)
    optimizer = training.make_optimizer(simulator.parameters(), config["lr_init"], device)   # fused, capturable, lr on the device
    step, lowest = 0, float("inf")
    if config.get("model_file") is not None:
        simulator, step, optimizer, lowest = load_model(simulator, device, config)
        training.set_lr(optimizer, training.learning_rate(step, config["lr_init"], config["lr_decay"],
                                                          config["lr_decay_steps"]))
    simulator.train()
    engines = {}                                   # particle count -> TrainEngine (captured step), at most two alive
    use_engine = config.get("train_engine", True) and train_engine.supports(simulator)
    loader = data_mod.get_data_loader_by_samples(
        str(Path(config["data_path"]) / "train.npz"), config["input_sequence_length"], config["batch_size"],
        shuffle=True, device=device, seed=seed, rank=rank, world=world)
    reducer = training.FlatGradAllReduce(simulator.parameters()) if world > 1 else None
    gen = torch.Generator(device=device).manual_seed(seed + 1000 * rank + step)
    save_dir = Path(config["model_path"]) / config["run_name"]
    t0, done = time.time(), False
    while not done:
        for b in loader:
            step += 1
            batch = {**b["input"], **b["output"]}
            n_glob = training.global_particle_count(batch["positions"].shape[0], device) if world > 1 else None
            eng = None
            if use_engine:
                n_loc = batch["positions"].shape[0]
                eng = engines.get(n_loc)
                if eng is None:
                    while len(engines) >= 2:
                        engines.pop(next(iter(engines)))
                    eng = engines[n_loc] = train_engine.TrainEngine(
                        simulator, optimizer, n_loc, device, w_pos=config["loss_weight_position"],
                        w_strain=config["loss_weight_strain"])
            loss = training.train_step(simulator, optimizer, batch, config["noise_std"],
                                       config["loss_weight_position"], config["loss_weight_strain"],
                                       reducer=reducer, n_global=n_glob, generator=gen, engine=eng)
            lr_new = training.learning_rate(step, config["lr_init"], config["lr_decay"], config["lr_decay_steps"])
            training.set_lr(optimizer, lr_new)
            if step % log_every == 0:
                if world > 1:
                    dist.all_reduce(loss)
                if rank == 0:
                    print(f"Step {step}: Total Loss = {loss.item():.6f}, lr = {lr_new:.3e}, "
                          f"{step / (time.time() - t0):.1f} steps/s")
            if step % config["nsave_steps"] == 0:
                if rank == 0 and (Path(config["data_path"]) / "valid.npz").exists():
                    eval_loss, _ = validate(simulator, metadata, config, device)
                    print(f"Mean loss on valid-set rollout prediction: {eval_loss}. Lowest so far {lowest}.")
                    if eval_loss < lowest:
                        lowest = eval_loss
                        save_checkpoint(simulator, optimizer, step, lowest, save_dir)
                if world > 1:
                    dist.barrier()
            if step >= config["ntraining_steps"]:
                done = True
                break
    return step, lowest


def main(argv: Optional[list] = None):
    ap = argparse.ArgumentParser(description="Single-Scale GNN Training (B200)")
    ap.add_argument("--config", type=str, default="sgnn/single_scale/config.yaml")
    ap.add_argument("--mode", type=str, choices=["train", "valid", "rollout"])
    ap.add_argument("--model_file", type=str)
    ap.add_argument("--log", type=str, choices=["True", "False"])
    args = ap.parse_args(argv)
    config = load_config(args.config)
    if args.mode is not None:
        config["mode"] = args.mode
    if args.model_file is not None:
        config["model_file"] = args.model_file
    if "LOCAL_RANK" in os.environ and int(os.environ.get("WORLD_SIZE", "1")) > 1:
        torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
        dist.init_process_group("nccl")
    device = torch.device("cuda", torch.cuda.current_device())
    with open(Path(config["data_path"]) / "metadata.json") as f:
        metadata = json.load(f)
    simulator = _get_simulator(metadata, config["noise_std"], config["noise_std"], device, config)
    if config["mode"] == "train":
        train(simulator, metadata, device, config)
    else:
        predict(simulator, metadata, device, config)
    if dist.is_initialized():
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
