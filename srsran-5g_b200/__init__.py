"""srsran-5g_b200: B200-native PUSCH FEC decode engine (rate dematch + HARQ combine -> LDPC -> CRC) behind a C ABI.

The directory name is not a Python identifier; load it with `importlib` (see __graft_entry__.load_package())."""
from . import build  # noqa: F401
from . import fec  # noqa: F401
