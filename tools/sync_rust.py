#!/usr/bin/env python
"""Regenerate rust/{build.rs,src/ffi.rs,src/dropin.rs} from the ```rust blocks of INTEGRATION.md (the single source)."""
import os, re
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
s = open(os.path.join(ROOT, "INTEGRATION.md")).read()
blocks = re.findall(r"```rust\n(.*?)```", s, flags=re.S)
hdr = ("// UNVERIFIED: written without a Rust toolchain (none exists in the build image); kept in step with INTEGRATION.md\n"
       "// by tools/sync_rust.py.  See INTEGRATION.md for where each piece goes in the reference tree.\n")
os.makedirs(os.path.join(ROOT, "rust", "src"), exist_ok=True)
open(os.path.join(ROOT, "rust", "build.rs"), "w").write(hdr + blocks[0])
open(os.path.join(ROOT, "rust", "src", "ffi.rs"), "w").write(hdr + blocks[1])
open(os.path.join(ROOT, "rust", "src", "dropin.rs"), "w").write(
    hdr + "// Replacements with the reference's own signatures: audio::log_mel_spectrogram, the LogitFilter / TokenExtractor /\n"
          "// SequenceRanker impls, Whisper::update_cache.\n" + "\n".join(blocks[2:]))
