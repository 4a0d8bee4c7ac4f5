from .comm import Comm, LocalComm, TorchComm, make_comm
from .reducer import Reducer, plan_buckets

__all__ = ["Comm", "LocalComm", "TorchComm", "make_comm", "Reducer", "plan_buckets"]
