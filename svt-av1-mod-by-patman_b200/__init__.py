"""B200-native open-loop motion estimation for SVT-AV1 (pyramid, pre-HME, HME L0-L2, full-pel
85-PU search) behind a C ABI (include/svtme_b200.h). `me` is the host-side convenience layer;
`abi` the ctypes mirror of the C structures."""
from . import abi  # noqa: F401
