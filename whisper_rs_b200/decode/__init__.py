"""Host-side mirror of reference src/decode/: LogitFilter / TokenExtractor / SequenceRanker with the
reference's names and call shapes, backed by the CUDA library (see filter.py, extractor.py, sequencer.py),
plus `FusedDecodeStep`, the fused replacement of src/decode/mod.rs:257-270, and `DecodeTask`, the loop that calls it
(src/decode/mod.rs:220-418)."""
from .core import DecodeOptions, DecodePrompt, Decoder, Projection, TokenExtractMode, build_decoder_cfg, max_initial_timestamp_index
from .filter import LogitFilter, SuppressBlank, SuppressTokens, TimestampTokens, build_filters
from .extractor import BeamSearchTokenExtractor, GreedyTokenExtractor, TokenExtractor, FusedDecodeStep
from .sequencer import MaximumLikelihoodRanker, SequenceRanker
from .task import DecodeOutput, DecodeTask

__all__ = [
    "Projection",
    "DecodeOptions", "DecodePrompt", "Decoder", "TokenExtractMode", "build_decoder_cfg", "max_initial_timestamp_index",
    "LogitFilter", "SuppressBlank", "SuppressTokens", "TimestampTokens", "build_filters",
    "BeamSearchTokenExtractor", "GreedyTokenExtractor", "TokenExtractor", "FusedDecodeStep",
    "MaximumLikelihoodRanker", "SequenceRanker", "DecodeOutput", "DecodeTask",
]
