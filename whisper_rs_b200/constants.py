"""Compile-time constants of the two hot paths.

Audio hyper-parameters mirror reference src/audio.rs:12-18.  Token-id sets mirror what
reference src/tokenize.rs:142-193 derives from tokenizer/gpt2.json (replayed offline by
tests/golden/make_golden.py -> tests/golden/tokenizer_constants.json; the ids are repeated here
so the product path never reads test fixtures).
"""
from dataclasses import dataclass, field
from typing import Tuple

# src/audio.rs:12-18
SAMPLE_RATE = 16000
N_FFT = 400
N_MELS = 80
HOP_LENGTH = 160
CHUNK_LENGTH = 30
N_SAMPLES = CHUNK_LENGTH * SAMPLE_RATE  # 480000
N_FRAMES = N_SAMPLES // HOP_LENGTH      # 3000
N_BINS = N_FFT // 2 + 1                 # 201

# src/tokenize.rs:215-307 applied to tokenizer/gpt2.json: sorted, NOT de-duplicated (96 entries, 84 unique)
NON_SPEECH_TOKENS_GPT2: Tuple[int, ...] = (
    1, 2, 7, 8, 9, 10, 14, 25, 26, 27, 28, 29, 31, 58, 59, 60, 61, 62, 63, 90, 91, 92, 93, 357, 366, 438, 532, 685,
    705, 796, 930, 1058, 1220, 1267, 1279, 1303, 1343, 1377, 1391, 1635, 1782, 1875, 2162, 2361, 2488, 3467, 4008,
    4211, 4600, 4808, 5299, 5855, 6329, 7203, 9609, 9959, 10563, 10786, 11420, 11709, 11907, 13163, 13697, 13700,
    14808, 15306, 16410, 16791, 17992, 17992, 17992, 17992, 17992, 17992, 17992, 19203, 19510, 20724, 20724, 20724,
    20724, 20724, 20724, 20724, 22305, 22935, 27007, 30109, 30420, 33409, 34949, 40283, 40493, 40549, 47282, 49146,
)


@dataclass(frozen=True)
class TokenIds:
    """The ids the logit filters / extractors need (pub fields of reference `Tokenizer`,
    src/tokenize.rs:113-131, plus the blank id used by src/decode/filter/suppress_blanks.rs:21-31)."""
    n_vocab: int
    eot: int
    sot: int
    translate: int
    transcribe: int
    startoflm: int
    startofprev: int
    no_speech: int
    no_timestamps: int
    ts_begin: int
    blank: int = 220  # encode(" ")[0] for the GPT-2 byte-level BPE
    non_speech_tokens: Tuple[int, ...] = field(default=NON_SPEECH_TOKENS_GPT2)

    def sequence_sot(self):
        # src/tokenize.rs:201-205 -- only [sot]
        return [self.sot]


# What the reference's Tokenizer::new_from_file("tokenizer/gpt2.json") yields (V = 51864, *.en models)
GPT2_EN = TokenIds(n_vocab=51864, eot=50256, sot=50257, translate=50357, transcribe=50358, startoflm=50359,
                   startofprev=50360, no_speech=50361, no_timestamps=50362, ts_begin=50363)

# BASELINE config 3: multilingual layout (everything after eot shifted by one; V = 51865).  The
# reference cannot build this tokenizer (src/tokenize.rs:144-148 asserts Transcribe/en), so these are
# parameters, not derived values.
MULTILINGUAL = TokenIds(n_vocab=51865, eot=50257, sot=50258, translate=50358, transcribe=50359, startoflm=50360,
                        startofprev=50361, no_speech=50362, no_timestamps=50363, ts_begin=50364)
