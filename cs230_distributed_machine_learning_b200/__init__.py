"""B200-native (candidate x CV-fold) sweep behind the distributed-ml API.

Drop-in for one path of sanjita2911/CS230-distributed-machine-learning: what
``MLTaskManager.train`` fans out (DistributedLibrary/src/distributed_ml/core.py:152)
and ``aws-prod/worker/worker.py:289`` executes one scikit-learn ``fit`` at a time.
"""
from .client import MLTaskManager  # noqa: F401
from .datasets import register_dataset, make_dataset  # noqa: F401

__all__ = ["MLTaskManager", "register_dataset", "make_dataset"]
