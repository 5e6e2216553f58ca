"""Entry points of the multi-scale simulator (reference sgnn/multi_scale/multi_scale_train.py:
load_config :36, predict :58, load_model :63, train :79, _get_simulator :290, main :358).  Same YAML
keys (sgnn/multi_scale/multi_scale_config.yaml), same flags (--config --mode --model_file --log), same
checkpoint files (model-best-{step:06}.pt / train_state-best-{step:06}.pt, model-final-* when no
validation ran) and the same rollout pickles.  Differences, all on the host side:

  * the data path is device-resident with one cached static graph per trajectory
    (static_graph_data_loader.py); a step does no host work and no H2D copy;
  * under torchrun training is data-parallel: ranks take disjoint slices of the seeded sample
    permutation ("shard by trajectory sample") and SUM-allreduce one flat gradient buffer
    (sgnn_b200/training.py); rank 0 validates and writes checkpoints;
  * validation compares `val/loss_total` (the reference reads a key its validator never returns --
    `rmse_total` / `eval_loss_mean`, :232 -- so its "best" is always the first validation);
  * a resumed run continues from the restored step (the reference restarts at 0, :124);
  * Weights & Biases logging is not wired (no network); `log` is accepted and ignored.
"""
import argparse
import json
import os
import time
from pathlib import Path
from typing import Optional

import torch
import torch.distributed as dist
import yaml

from .. import training
from .multi_scale_evaluate import validate_multi_scale_simulator
from .multi_scale_inference import run_inference
from .multi_scale_simulator import MultiScaleSimulator
from .static_graph_data_loader import get_multi_scale_data_loader_by_samples


def load_config(config_path):
    """YAML -> dict; relative paths are tried against the working directory, then against this
    package (reference :36-52)."""
    config_file = Path(config_path)
    if not config_file.is_absolute():
        if config_file.exists():
            config_file = config_file.resolve()
        elif (Path(__file__).parent / config_file).exists():
            config_file = (Path(__file__).parent / config_file).resolve()
        else:
            raise FileNotFoundError(f"Config file not found: {config_path}")
    with open(config_file) as f:
        config = yaml.safe_load(f)
    config.setdefault("train_state_file", None)
    config.setdefault("inference_mode", "autoregressive")
    config.setdefault("loss_weight_position", 1.0)
    config.setdefault("loss_weight_strain", 1.0)
    return config


def predict(simulator: MultiScaleSimulator, metadata: dict, device, config: dict):
    """Delegates to multi_scale_inference.run_inference (reference :58-60)."""
    return run_inference(simulator, metadata, config, device)


def load_model(simulator, device, config: dict):
    """(simulator, step, optimizer, lowest_eval_loss) from model_file + train_state_file under
    model_path/run_name (reference :63-76, utils/checkpoint_utils.py:13-44)."""
    d = Path(config["model_path"]) / config["run_name"]
    mp, sp = d / config["model_file"], d / str(config["train_state_file"])
    if not (mp.exists() and sp.exists()):
        raise FileNotFoundError(f"Specified model_file {mp} and train_state_file {sp} not found.")
    simulator.load(str(mp))
    simulator.to(device)
    state = torch.load(str(sp), map_location="cpu")
    optimizer = training.make_optimizer(simulator.parameters(), 1e-3, device)
    optimizer.load_state_dict(state["optimizer_state"])
    gts = state["global_train_state"]
    return simulator, int(gts["step"]), optimizer, float(gts.get("lowest_eval_loss", float("inf")))


def _save(simulator, optimizer, step, lowest, save_dir: Path, tag: str):
    save_dir.mkdir(parents=True, exist_ok=True)
    simulator.save(str(save_dir / f"model-{tag}-{step:06}.pt"))
    gts = {"step": step} if tag == "final" else {"step": step, "lowest_eval_loss": lowest}
    torch.save({"optimizer_state": optimizer.state_dict(), "global_train_state": gts},
               str(save_dir / f"train_state-{tag}-{step:06}.pt"))


def train(simulator: MultiScaleSimulator, metadata: dict, device, config: dict, log_every: int = 10, seed: int = 0):
    """Training loop (reference :79-287).  Returns (last step, lowest validation loss)."""
    rank = dist.get_rank() if dist.is_initialized() else 0
    world = dist.get_world_size() if dist.is_initialized() else 1
    simulator.to(device)
    optimizer = training.make_optimizer(simulator.parameters(), config["lr_init"], device)
    step, lowest = 0, float("inf")
    if config.get("model_file"):
        try:
            simulator, step, optimizer, lowest = load_model(simulator, device, config)
            training.set_lr(optimizer, training.learning_rate(step, config["lr_init"], config["lr_decay"],
                                                              config["lr_decay_steps"]))
            if rank == 0:
                print(f"Resumed training from step {step}")
        except FileNotFoundError as e:           # reference :107-112: start fresh
            if rank == 0:
                print(f"Resume requested but checkpoint not found: {e}. Starting fresh.")
    simulator.train()
    simulator.to(device)
    loader = get_multi_scale_data_loader_by_samples(
        path=str(Path(config["data_path"]) / "train.npz"), input_length_sequence=config["input_sequence_length"],
        batch_size=config["batch_size"], shuffle=True, num_scales=config["num_scales"],
        window_size=config["window_size"], radius_multiplier=config["radius_multiplier"], device=device, seed=seed,
        rank=rank, world=world)
    reducer = training.FlatGradAllReduce(simulator.parameters()) if world > 1 else None
    gen = torch.Generator(device=device).manual_seed(seed + 1000 * rank + step)
    save_dir = Path(config["model_path"]) / config["run_name"]
    valid_path = Path(config["data_path"]) / "valid.npz"
    t0, done = time.time(), False
    while not done:
        for b in loader:
            step += 1
            simulator.set_static_graph(b["graph"])          # the batch's graph = its first sample's (:143, collate quirk)
            batch = {**b["input"], **b["output"]}
            n_glob = training.global_particle_count(batch["positions"].shape[0], device) if world > 1 else None
            loss = training.train_step(simulator, optimizer, batch, config["noise_std"],
                                       config["loss_weight_position"], config["loss_weight_strain"],
                                       reducer=reducer, n_global=n_glob, generator=gen)
            lr_new = training.learning_rate(step, config["lr_init"], config["lr_decay"], config["lr_decay_steps"])
            training.set_lr(optimizer, lr_new)
            if step % log_every == 0:
                if world > 1:
                    dist.all_reduce(loss)
                if rank == 0:
                    print(f"Step {step}: Total Loss = {loss.item():.6f}, lr = {lr_new:.3e}, "
                          f"{step / (time.time() - t0):.1f} steps/s")
            if step % config["nsave_steps"] == 0:
                if rank == 0 and valid_path.exists():
                    metrics = validate_multi_scale_simulator(
                        simulator=simulator, data_path=str(valid_path), metadata=metadata, device=device,
                        input_sequence_length=config["input_sequence_length"], inference_mode=config["inference_mode"],
                        num_scales=config["num_scales"], window_size=config["window_size"],
                        radius_multiplier=config["radius_multiplier"], verbose=False)
                    eval_loss = metrics["val/loss_total"]
                    print(f"Mean loss on valid-set rollout prediction: {eval_loss}. Current lowest eval loss is {lowest}.")
                    if eval_loss < lowest:
                        lowest = eval_loss
                        _save(simulator, optimizer, step, lowest, save_dir, "best")
                    simulator.train()
                if world > 1:
                    dist.barrier()
            if step >= config["ntraining_steps"]:
                done = True
                break
    if lowest == float("inf") and rank == 0:        # no validation ran: keep the final weights (reference :262-271)
        _save(simulator, optimizer, step, lowest, save_dir, "final")
    return step, lowest


def _get_simulator(metadata: dict, acc_noise_std: float, vel_noise_std: float, device, config: dict
                   ) -> MultiScaleSimulator:
    """Reference :290-355: std = sqrt(std^2 + noise_std^2); nnode_in = (S-1)*dim + 1 (+ embedding with several
    particle types); nedge_in = dim + 1; nedge_out = latent = hidden_dim; nmlp_layers = 2."""
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
    return MultiScaleSimulator(
        kinematic_dimensions=config["dim"], nnode_in=nnode_in, nedge_in=config["dim"] + 1,
        nedge_out=config["hidden_dim"], latent_dim=config["hidden_dim"], nmessage_passing_steps=config["layers"],
        nmlp_layers=2, normalization_stats=stats, nparticle_types=ntypes,
        particle_type_embedding_size=config["particle_type_embedding_size"], num_scales=config["num_scales"],
        window_size=config["window_size"], radius_multiplier=config["radius_multiplier"], device=device)


def main(argv: Optional[list] = None):
    ap = argparse.ArgumentParser(description="Multi-Scale GNN Training (B200)")
    ap.add_argument("--config", type=str, default="sgnn/multi_scale/multi_scale_config.yaml")
    ap.add_argument("--mode", type=str, choices=["train", "valid", "rollout"])
    ap.add_argument("--model_file", type=str)
    ap.add_argument("--log", type=str, choices=["True", "False"])
    args = ap.parse_args(argv)
    config = load_config(args.config)
    if args.mode is not None:
        config["mode"] = args.mode
    if args.model_file is not None:
        config["model_file"] = args.model_file
    if args.log is not None:
        config["log"] = args.log == "True"
    if "LOCAL_RANK" in os.environ and int(os.environ.get("WORLD_SIZE", "1")) > 1:
        torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
        dist.init_process_group("nccl")
    device = torch.device("cuda", torch.cuda.current_device())
    with open(Path(config["data_path"]) / "metadata.json") as f:
        metadata = json.load(f)
    simulator = _get_simulator(metadata, config["noise_std"], config["noise_std"], device, config)
    simulator.to(device)
    if config["mode"] == "train":
        train(simulator, metadata, device, config)
    elif config["mode"] == "valid":
        if config.get("model_file"):
            model_path = Path(config["model_path"]) / config["run_name"] / config["model_file"]
            if not model_path.exists():
                print(f"Model file not found: {model_path}")
                return None
            simulator.load(str(model_path))
        metrics = validate_multi_scale_simulator(
            simulator=simulator, data_path=str(Path(config["data_path"]) / "valid.npz"), metadata=metadata,
            device=device, input_sequence_length=config["input_sequence_length"],
            inference_mode=config["inference_mode"], num_scales=config["num_scales"],
            window_size=config["window_size"], radius_multiplier=config["radius_multiplier"])
        for key, value in metrics.items():
            print(f"   {key}: {value:.6f}")
        return metrics
    elif config["mode"] == "rollout":
        predict(simulator, metadata, device, config)
    if dist.is_initialized():
        dist.destroy_process_group()
    return None


if __name__ == "__main__":
    main()
