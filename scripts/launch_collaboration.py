#!/usr/bin/env python
"""One-box fleet launcher (A7 / S8: what albert/AWS_runner.ipynb and sahajbert/contributor_notebook.ipynb do with
boto3 spot fleets and Colab cells): 1 coordinator + N trainer peers (one per GPU) + M auxiliary peers, a respawn
loop that keeps the fleet at size (spot-preemption behaviour: a killed peer comes back, resumes from its newest
`checkpoint*` and immediately overwrites it with the collaboration's state), and optional fault injection
(`--kill_every S` kills a random trainer every S seconds; `--pause_every S` SIGSTOPs one for `--pause_for`).
Bandwidth shaping (wondershaper in the reference) is meaningless on NVLink and is not emulated.
`--fleet_report FILE` turns a run into the elasticity benchmark of BASELINE.json config 3: collaboration-wide samples/s
over the window after `--measure_after` seconds (from the coordinator's view of the global step), faults injected,
rounds that failed / ran solo, state downloads and the time a respawned peer needs to contribute again.

    python scripts/launch_collaboration.py --experiment_prefix demo --trainers 8 --aux 1 --synthetic_data \
        --target_batch_size 4096 --per_device_train_batch_size 64 --run_for 600
"""
import argparse
import os
import random
import re
import signal
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def spawn(module, args, env_extra=None, log=None):
    env = dict(os.environ)
    env["PYTHONPATH"] = ROOT + os.pathsep + env.get("PYTHONPATH", "")
    env.update(env_extra or {})
    out = open(log, "a") if log else None
    return subprocess.Popen([sys.executable, "-u", "-m", module, *args], env=env, stdout=out, stderr=subprocess.STDOUT,
                            cwd=ROOT, start_new_session=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--experiment_prefix", default="demo")
    ap.add_argument("--family", default="albert", choices=["albert", "sahajbert"])
    ap.add_argument("--trainers", type=int, default=1)
    ap.add_argument("--aux", type=int, default=0)
    ap.add_argument("--clients", type=int, default=0, help="how many of the trainers run with --client_mode")
    ap.add_argument("--target_batch_size", type=int, default=4096)
    ap.add_argument("--per_device_train_batch_size", type=int, default=64)
    ap.add_argument("--gradient_accumulation_steps", type=int, default=1)
    ap.add_argument("--compression", default="FLOAT16")
    ap.add_argument("--synthetic_data", action="store_true")
    ap.add_argument("--config_path", default="")
    ap.add_argument("--uplinks", default="",
                    help="comma-separated Mbit/s, cycled over the trainers (the reference's AWS fleet: 200,200,200,200,"
                         "100,100,100,100,100,100,100,100,50,50,50,50): sets --bandwidth (load balancing) and "
                         "--emulated_uplink_mbps (pacing of the socket all-reduce) of each trainer; --aux_uplink for aux peers")
    ap.add_argument("--aux_uplink", type=float, default=0.0)
    ap.add_argument("--rtt_ms", type=float, default=0.0,
                    help="every RPC of every process waits this long first (DEDLOC_EMULATED_RTT_MS): internet round trips")
    ap.add_argument("--run_for", type=float, default=0, help="seconds; 0 = until interrupted")
    ap.add_argument("--kill_every", type=float, default=0)
    ap.add_argument("--pause_every", type=float, default=0)
    ap.add_argument("--pause_for", type=float, default=10)
    ap.add_argument("--log_dir", default="fleet_logs")
    ap.add_argument("--device", default=None, help="cpu to run the fleet without GPUs")
    ap.add_argument("--fleet_report", default="", help="write the elasticity report here when the run ends")
    ap.add_argument("--measure_after", type=float, default=60.0, help="seconds of start-up excluded from the report")
    args, passthrough = ap.parse_known_args()
    if args.rtt_ms > 0:
        os.environ["DEDLOC_EMULATED_RTT_MS"] = str(args.rtt_ms)      # inherited by everything spawn() starts
    os.makedirs(args.log_dir, exist_ok=True)
    pkg = f"dedloc_b200.trainers.{args.family}"

    coord_log = os.path.join(args.log_dir, "coordinator.log")
    coord = spawn(f"{pkg}.run_first_peer", ["--experiment_prefix", args.experiment_prefix, "--address", "127.0.0.1",
                                            "--dht_listen_on", "127.0.0.1:*", "--refresh_period", "5",
                                            "--target_batch_size", str(args.target_batch_size)], log=coord_log)
    address = None
    for _ in range(600):
        time.sleep(0.2)
        if os.path.exists(coord_log):
            m = re.search(r"Running DHT root at ([\d.]+:\d+)", open(coord_log).read())
            if m:
                address = m.group(1)
                break
    assert address, f"coordinator did not start; see {coord_log}"
    print(f"coordinator up at {address}", flush=True)

    def trainer_args(i):
        a = ["--experiment_prefix", args.experiment_prefix, "--initial_peers", address,
             "--target_batch_size", str(args.target_batch_size), "--compression", args.compression,
             "--per_device_train_batch_size", str(args.per_device_train_batch_size),
             "--gradient_accumulation_steps", str(args.gradient_accumulation_steps),
             "--output_dir", os.path.join(args.log_dir, f"outputs_{i}"), "--dht_listen_on", "127.0.0.1:*",
             "--listen_on", "127.0.0.1:*", "--min_refresh_period", "0.05", "--default_refresh_period", "0.2",
             "--max_refresh_period", "1.0", "--target_group_size", str(max(args.trainers + args.aux, 2))]
        if args.synthetic_data:
            a.append("--synthetic_data")
        if args.config_path:
            a += ["--config_path", args.config_path]
        if i < args.clients:
            a.append("--client_mode")
        if args.device:
            a += ["--device", args.device]
        uplinks = [float(x) for x in args.uplinks.split(",") if x]
        rate = args.aux_uplink if i >= args.trainers else (uplinks[i % len(uplinks)] if uplinks else 0.0)
        if rate > 0:
            a += ["--bandwidth", str(rate), "--emulated_uplink_mbps", str(rate)]
        return a + passthrough

    fleet = {}

    def ensure_fleet():
        for i in range(args.trainers):
            key = ("trainer", i)
            if key not in fleet or fleet[key].poll() is not None:
                env = {} if args.device == "cpu" else {"CUDA_VISIBLE_DEVICES": str(i)}
                fleet[key] = spawn(f"{pkg}.run_trainer", trainer_args(i), env, os.path.join(args.log_dir, f"trainer_{i}.log"))
                print(f"(re)started trainer {i} pid {fleet[key].pid}", flush=True)
                if key in started_once:
                    events["respawns"].append({"trainer": i, "t": time.time() - t0})
                started_once.add(key)
        for j in range(args.aux if args.family == "albert" else 0):
            key = ("aux", j)
            if key not in fleet or fleet[key].poll() is not None:
                a = [x for x in trainer_args(args.trainers + j) if x != "--synthetic_data"] + ["--device", "cpu"]
                fleet[key] = spawn(f"{pkg}.run_aux", a, {"CUDA_VISIBLE_DEVICES": ""}, os.path.join(args.log_dir, f"aux_{j}.log"))
                print(f"(re)started aux {j} pid {fleet[key].pid}", flush=True)

    t0 = last_kill = last_pause = time.time()
    paused = None
    events = {"kills": [], "pauses": [], "respawns": []}
    started_once = set()
    try:
        while not args.run_for or time.time() - t0 < args.run_for:
            ensure_fleet()
            now = time.time()
            trainers = [k for k in fleet if k[0] == "trainer" and fleet[k].poll() is None]
            if args.kill_every and now - last_kill > args.kill_every and trainers:
                victim = random.choice(trainers)
                os.killpg(fleet[victim].pid, signal.SIGKILL)     # exact process group we started
                print(f"fault injection: killed {victim}", flush=True)
                events["kills"].append({"trainer": victim[1], "t": now - t0})
                last_kill = now
            if args.pause_every and paused is None and now - last_pause > args.pause_every and trainers:
                victim = random.choice(trainers)
                os.killpg(fleet[victim].pid, signal.SIGSTOP)
                paused = (victim, now)
                print(f"fault injection: paused {victim}", flush=True)
                events["pauses"].append({"trainer": victim[1], "t": now - t0, "for": args.pause_for})
            if paused is not None and now - paused[1] > args.pause_for:
                os.killpg(fleet[paused[0]].pid, signal.SIGCONT)
                print(f"fault injection: resumed {paused[0]}", flush=True)
                paused, last_pause = None, now
            time.sleep(1.0)
    except KeyboardInterrupt:
        pass
    finally:
        for p in list(fleet.values()) + [coord]:
            if p.poll() is None:
                try:
                    os.killpg(p.pid, signal.SIGTERM)
                except ProcessLookupError:
                    pass
        print("fleet stopped", flush=True)
        if args.fleet_report:
            write_report(args, events, t0, coord_log)


def _stamp(line: str):
    """Seconds since the epoch of a log line that starts with a time stamp, or None."""
    import datetime

    for fmt in ("%m/%d/%Y %H:%M:%S", "%Y-%m-%d %H:%M:%S"):      # trainers / coordinator
        try:
            return datetime.datetime.strptime(line[:19], fmt).timestamp()
        except ValueError:
            continue
    return None


def write_report(args, events, t0, coord_log):
    import json

    # the coordinator logs "Step #N loss = L" (run_first_peer.py) whenever the collaboration's step advanced
    steps = []
    for line in open(coord_log, errors="replace"):
        m = re.search(r"Step #(\d+)", line)
        ts = _stamp(line)
        if m and ts is not None:
            steps.append((ts - t0, int(m.group(1))))
    window = [(t, s) for t, s in steps if t >= args.measure_after]
    report = {"family": args.family, "trainers": args.trainers, "aux": args.aux, "clients": args.clients,
              "compression": args.compression, "target_batch_size": args.target_batch_size,
              "micro_batch": args.per_device_train_batch_size, "run_for_s": args.run_for,
              "measure_after_s": args.measure_after, "faults": events}
    if len(window) >= 2:
        (ta, sa), (tb, sb) = window[0], window[-1]
        report.update(window_s=round(tb - ta, 1), collaborative_steps_in_window=sb - sa,
                      samples_per_second=round((sb - sa) * args.target_batch_size / max(tb - ta, 1e-9), 1))
    # the same figure from the trainers' own per-round lines (1 s stamps instead of the coordinator's polling period):
    # a collaborative step counts when the FIRST trainer reports it
    first_seen = {}
    for i in range(args.trainers):
        path = os.path.join(args.log_dir, f"trainer_{i}.log")
        if not os.path.exists(path):
            continue
        for line in open(path, errors="replace"):
            m = re.search(r"collaborative step (\d+) done", line)
            ts = _stamp(line)
            if m and ts is not None:
                step = int(m.group(1))
                first_seen[step] = min(first_seen.get(step, ts), ts)
    in_window = sorted((ts - t0, step) for step, ts in first_seen.items() if ts - t0 >= args.measure_after)
    if len(in_window) >= 2:
        (ta, sa), (tb, sb) = in_window[0], in_window[-1]
        report.update(trainer_log_window_s=round(tb - ta, 1), trainer_log_steps=sb - sa,
                      samples_per_second_from_trainer_logs=round((sb - sa) * args.target_batch_size / max(tb - ta, 1e-9), 1))
    per_trainer = {}
    rejoin = []
    for i in range(args.trainers):
        path = os.path.join(args.log_dir, f"trainer_{i}.log")
        if not os.path.exists(path):
            continue
        text = open(path, errors="replace").read()
        per_trainer[i] = {"failed_rounds": text.count("Skipped averaging"),
                          "state_downloads": len(re.findall(r"downloading state|Loading state from peers", text)),
                          "torn_rounds": text.count("must resynchronise after a torn round"),
                          "solo_or_local_steps_logged": text.count("averaging round failed"),
                          "collaborative_steps_logged": len(re.findall(r"-\s+Step \d+$", text, flags=re.M))}
        # time from a respawn to the first "Step N" line of the new process
        stamps = [(_stamp(l), l) for l in text.splitlines()]
        for ev in events["respawns"]:
            if ev["trainer"] != i:
                continue
            after = [ts - t0 for ts, l in stamps if ts is not None and ts - t0 > ev["t"] and re.search(r"-\s+Step \d+$", l)]
            if after:
                rejoin.append(round(after[0] - ev["t"], 1))
    report["per_trainer"] = per_trainer
    report["seconds_from_respawn_to_first_collaborative_step"] = rejoin
    json.dump(report, open(args.fleet_report, "w"), indent=1)
    print(json.dumps({k: v for k, v in report.items() if k != "per_trainer"}), flush=True)


if __name__ == "__main__":
    main()
