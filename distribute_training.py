#!/usr/bin/env python
"""Drop-in entry point under the reference's file name: the command lines from its README keep working.

    python distribute_training.py --job_name=ps     --task_id=0 --ps_hosts=... --worker_hosts=...
    python distribute_training.py --job_name=worker --task_id=0 --sync_replicas=True ...

Everything is implemented in ``tensorflow_train_distributed_b200.train`` (flags: ``config.py``).
"""
import sys

from tensorflow_train_distributed_b200.train import main

if __name__ == "__main__":
    sys.exit(main())
