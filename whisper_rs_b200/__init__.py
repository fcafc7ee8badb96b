"""B200-native hot paths of laptou/whisper-rs: the log-mel front-end (audio) and the per-step logit
processing (decode), behind the C ABI of include/wrs.h.  Importing this package never touches the GPU;
the first compute call loads whisper_rs_b200/lib/libwrs_b200.so and fails loudly if it is absent."""
from . import constants  # noqa: F401

__version__ = "0.1.0"
