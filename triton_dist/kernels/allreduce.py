"""AllReduce method enumeration (same names/values as the reference's
``python/triton_dist/kernels/allreduce.py:28-66``)."""
from enum import IntEnum


class AllReduceMethod(IntEnum):
    Unknown = 0
    OneShot = 1
    TwoShot = 2
    DoubleTree = 3
    OneShot_TMA = 4
    OneShot_Multimem = 5
    TwoShot_Multimem = 6
    TwoShot_Multimem_ST = 7
    OneShot_LL = 8
    OneShot_Multimem_LL = 9
    AllReduceEnumMax = 10


_NAMES = {
    "double_tree": AllReduceMethod.DoubleTree,
    "one_shot": AllReduceMethod.OneShot,
    "two_shot": AllReduceMethod.TwoShot,
    "one_shot_tma": AllReduceMethod.OneShot_TMA,
    "one_shot_multimem": AllReduceMethod.OneShot_Multimem,
    "two_shot_multimem": AllReduceMethod.TwoShot_Multimem,
    "two_shot_multimem_st": AllReduceMethod.TwoShot_Multimem_ST,
}


def to_allreduce_method(method: str) -> AllReduceMethod:
    try:
        return _NAMES[method]
    except KeyError:
        raise ValueError(f"Invalid method name {method}. Supported methods: {list(_NAMES)}") from None


def get_allreduce_methods():
    return list(_NAMES)
