"""sdrx_b200 -- B200-native per-channel DSP chain of johanhedin/sdrx (channelizer + AM demod + squelch).

The compute path is the CUDA library ``libsdrx_b200.so`` (sm_100a) behind the C ABI of
``include/sdrx_b200.h``; this package is its host-side mirror.  There is no CPU fallback.
"""
from ._lib import FMT_F32, FMT_S16, FMT_U8, SdrxError  # noqa: F401
from .config import ChannelPlan, rate_config  # noqa: F401
from .engine import BlockOut, Engine, PinnedBuffer  # noqa: F401
from .source import Carrier, FileSource, SyntheticSource  # noqa: F401
from .airspy_fe import AirspyFrontEnd  # noqa: F401,E402
