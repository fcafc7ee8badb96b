"""CPU oracle for the two hot paths of laptou/whisper-rs.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline / `--impl reference` legs may import
this package; it is the checker, never the product path (the product fails loudly when the CUDA
library is missing).

What it is: a restatement, op for op, of the reference's ATen call sequence on CPU.  The reference is
Rust over `tch` 0.10 (git laptou/tch-rs@0d0d6dbf, libtorch 1.13.0; Cargo.lock:1225-1227,
.vscode/settings.json:7) and contains no arithmetic of its own on these paths -- every number comes
out of libtorch.  No Rust toolchain exists in the build image, the dependencies are un-vendored and
there is no network, so the reference itself cannot be built (`oracle/_ref` does not exist; DESIGN.md
§Oracle).  The same ATen operators are issued here through Python torch 2.11 (CPU).

Parity pinning status:
  * mel filterbank          PINNED  by the reference's own KAT (src/audio.rs:266-279,
                                    test/data/mel-filter-8x8.csv; tests/golden/reference_kat.json)
  * log-mel spectrogram     PARTLY  pinned: the golden (test/data/mel-spectrogram.csv) pins shape,
                                    floor (max-8) and (x+4)/4 scaling; its input WAV is absent from the
                                    reference repo, so sample values are pinned only by the op sequence
  * filters / beam / ranker UNPINNED by the reference (it has no test for src/decode/**): "parity
                                    unpinned" -- pinned only by this restatement and the committed
                                    regression vectors in tests/golden/oracle_vectors.npz
  * greedy                  UNPINNED and not runnable in the reference (todo!() at
                                    src/decode/mod.rs:134); restates the intent of extractor/greedy.rs
"""
