"""Boundary-condition tags and basic types of the reference.

Mirrors /root/reference/include/BCtypes.hpp:6-16 (enum BCTag) and
/root/reference/include/Types.hpp:20-33 (index_type=int32, real_type=double).
"""
import numpy as np

index_type = np.int32
real_type = np.float64


class BCTag:
    In = 1
    Out = 2
    Wall = 3
    Far = 4
    Cyl = 5
    Dirichlet = 6
    Neuman = 7
    Slip = 8


NUM_BC_TAGS = 9  # tags are 1..8; slot 0 unused (interior faces carry 0)
