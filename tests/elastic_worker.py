"""Helper process for tests/test_gpu_multi.py::test_late_joiner_and_peer_death (not collected by pytest).

role=first : GPU 0. Starts the DHT, trains until `--until` collaborative steps, records a digest of the fp32 master
             parameters after every step, survives the other peer vanishing.
role=second: GPU 1. Joins late (after the first peer reached step 2), must download the state from the first peer,
             takes part in joint rounds until `--until`, writes {step: digest} and dies WITHOUT any shutdown.
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch


def digest(t: torch.Tensor) -> str:
    t = t.double()
    return f"{float(t.sum()):.12e}/{float(t.abs().sum()):.12e}/{float((t * t).sum()):.12e}"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--role", required=True)
    ap.add_argument("--dir", required=True)
    ap.add_argument("--until", type=int, required=True)
    ap.add_argument("--device", default="cuda")
    args = ap.parse_args()
    from dedloc_b200.dht import DHT
    from dedloc_b200.models.albert import AlbertConfig
    from dedloc_b200.trainers.pretrain import PeerConfig, SyntheticMLMBatches, build_albert_peer

    first = args.role == "first"
    dev = torch.device("cuda", 0 if first else 1) if args.device == "cuda" else torch.device("cpu")
    if dev.type == "cuda":
        torch.cuda.set_device(dev)
    port_file = os.path.join(args.dir, "port")
    prog_file = os.path.join(args.dir, "first_step")
    if first:
        dht = DHT(start=True)
        open(port_file, "w").write(str(dht.port))
    else:
        while not os.path.exists(prog_file) or int(open(prog_file).read() or 0) < 2:
            time.sleep(0.05)
        dht = DHT(start=True, initial_peers=[f"127.0.0.1:{open(port_file).read().strip()}"])
    model_cfg = AlbertConfig.small()
    cfg = PeerConfig(prefix="elastic", target_batch_size=64, micro_batch=8, seq_len=128, compression="FLOAT16",
                     target_group_size=2, averaging_expiration=1.0, averaging_timeout=6.0, metadata_expiration=2.5,
                     min_refresh_period=0.05, default_refresh_period=0.1, max_refresh_period=0.5)
    trainer = build_albert_peer(dht, cfg, model_cfg, device=dev, seed=0 if first else 1)
    data = SyntheticMLMBatches(model_cfg.vocab_size, 128, 8, seed=11 + (0 if first else 1))
    batches = [data.to_device(b, dev) for b in data.batches]
    collab = trainer.collab
    digests = {}
    t0 = time.time()
    i = 0
    joined_at = None
    # first : runs until the second peer has vanished and `--until` more collaborative steps have happened after that
    # second: runs `--until` collaborative steps after it joined, then dies
    other_json = os.path.join(args.dir, "second.json")
    death_step = None
    while time.time() - t0 < 200:
        if first:
            if death_step is None and os.path.exists(other_json):
                death_step = collab.local_step
            if death_step is not None and collab.local_step >= death_step + args.until:
                break
            time.sleep(0.02)          # GPU peers step every ~15 ms: do not run hundreds of steps while the other starts
        elif joined_at is not None and collab.local_step >= joined_at + args.until:
            break
        stepped = trainer.micro_step(batches[i % len(batches)])
        i += 1
        if not first and joined_at is None and collab.local_step > 0:
            joined_at = collab.local_step
        if stepped:
            if dev.type == "cuda":
                torch.cuda.synchronize()
                digests[collab.local_step] = digest(trainer.opt.buffers.p32)
            else:
                digests[collab.local_step] = digest(torch.cat([p.detach().float().reshape(-1) for p in trainer.model.parameters()]))
            if first:
                open(prog_file, "w").write(str(collab.local_step))
    out = {"role": args.role, "local_step": collab.local_step, "stats": {k: v for k, v in collab.stats.items() if isinstance(v, (int, float))},
           "digests": digests, "joined_at": joined_at, "micro_steps": i, "death_step": death_step}
    json.dump(out, open(os.path.join(args.dir, f"{args.role}.json"), "w"))
    print(json.dumps({k: v for k, v in out.items() if k != "digests"}), flush=True)
    if first:
        collab.shutdown()
        dht.shutdown()
    else:
        os._exit(0)          # a preempted peer: no goodbye, its DHT records simply expire


if __name__ == "__main__":
    main()
