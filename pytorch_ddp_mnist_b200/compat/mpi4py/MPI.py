"""Singleton-world `mpi4py.MPI` subset; see the package docstring."""
import numpy as _np

MAX, MIN, SUM = "MAX", "MIN", "SUM"


class _Comm:
    def Get_rank(self) -> int:
        return 0

    def Get_size(self) -> int:
        return 1

    rank = property(Get_rank)
    size = property(Get_size)

    def Barrier(self) -> None:
        return None

    barrier = Barrier

    def Reduce(self, sendbuf, recvbuf, op=SUM, root: int = 0) -> None:
        # one rank: every reduction is the identity (`mp:264`, dead code in the reference)
        _np.copyto(_np.asarray(recvbuf), _np.asarray(sendbuf))

    def Allreduce(self, sendbuf, recvbuf, op=SUM) -> None:
        self.Reduce(sendbuf, recvbuf, op)

    def bcast(self, obj, root: int = 0):
        return obj


COMM_WORLD = _Comm()
COMM_SELF = _Comm()


def Is_initialized() -> bool:
    # False keeps the reference's fallbacks on the env-variable branch (`mp:232-238`)
    return False


def Is_finalized() -> bool:
    return False
