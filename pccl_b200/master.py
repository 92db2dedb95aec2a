"""`pccl_master --listen-address ip:port` console entry (reference: python/framework/pccl/master.py)."""
import argparse
import signal
import time

from ._pccl import MasterNode


def main() -> None:
    ap = argparse.ArgumentParser(description="CCoIP master node")
    ap.add_argument("--listen-address", default="0.0.0.0:48148")
    args = ap.parse_args()
    node = MasterNode(args.listen_address)
    node.run()
    print(f"master listening on port {node.port}", flush=True)
    stop = []
    signal.signal(signal.SIGINT, lambda *_: stop.append(1))
    signal.signal(signal.SIGTERM, lambda *_: stop.append(1))
    while not stop:
        time.sleep(0.2)
    node.interrupt()


if __name__ == "__main__":
    main()
