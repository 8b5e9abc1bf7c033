"""tensorflow_train_distributed_b200 — a Blackwell (B200, sm_100a) native parameter-server training framework
with the capabilities of boyuanf/tensorflow_train_distributed's ``distribute_training.py``.

Layers:

* ``config`` / ``cluster``        flags, ClusterSpec-style role map          (reference :22-36, :168-183)
* ``parallel``                    sharding plans, parameter store, sync/async replica updaters, CPU/gloo ps
* ``ops``                         ctypes bindings of the hand-written sm_100a kernels (``csrc/*.cu``)
* ``models``                      the reference MLP, the wide MLP and LeNet-5 on top of those kernels
* ``session``                     MonitoredTrainingSession-style driver with hooks (reference :196-238)
* ``io``                          TF1 tensor-bundle checkpoints and events files (C++ in ``csrc/io``)
"""

__version__ = "0.1.0"
