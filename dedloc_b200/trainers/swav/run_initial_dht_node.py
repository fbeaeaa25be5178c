#!/usr/bin/env python
"""SwAV DHT bootstrap node (W1; behaviour of /root/reference/swav/run_initial_dht_node.py:1-40): start a
listening DHT node, print its address for `+config.OPTIMIZER.dht_initial_peers`, and keep the routing table warm by
looking up a random key every `refresh_period` seconds."""
import argparse
import logging
import time
import uuid

from ...dht import DHT

logger = logging.getLogger(__name__)


def main(argv=None):
    parser = argparse.ArgumentParser()
    parser.add_argument("--address", type=str, required=False, help="this machine's network address")
    parser.add_argument("--listen_on", type=str, default="0.0.0.0:*", required=False, help="interface:port to bind")
    parser.add_argument("--refresh_period", type=float, default=30, required=False,
                        help="fetch a random key from the DHT this often (seconds)")
    parser.add_argument("--max_iterations", type=int, default=0, help="stop after this many refreshes (0 = forever)")
    args = parser.parse_args(argv)
    logging.basicConfig(level=logging.INFO, format="%(asctime)s [%(levelname)s] %(message)s")
    listen_on = args.listen_on.replace("[::]", "0.0.0.0")
    dht = DHT(start=True, listen_on=listen_on, endpoint=f"{args.address}:*" if args.address else None)
    logger.info(f"Running DHT root at {args.address or '127.0.0.1'}:{dht.port}")
    print(f"INITIAL_DHT_ADDRESS={args.address or '127.0.0.1'} INITIAL_DHT_PORT={dht.port}", flush=True)
    n = 0
    while True:
        dht.get(uuid.uuid4().bytes, latest=True)
        n += 1
        if args.max_iterations and n >= args.max_iterations:
            break
        time.sleep(args.refresh_period)
    dht.shutdown()
    return dht


if __name__ == "__main__":
    main()
