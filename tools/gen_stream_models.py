#!/usr/bin/env python
"""Generate build/driver/StreamModels.h = the reference's src/StreamModels.h + the B200 branch.

    python tools/gen_stream_models.py <reference StreamModels.h> <output>

Upstream, registering a model is three edits (cmake/register_models.cmake, CMakeLists.txt:161-175,
src/StreamModels.h:4-34 and :36-99).  The reference tree is read-only here, so the two
StreamModels.h edits are applied to a generated copy at build time; nothing of the reference is
committed to this repository.  The output is the input text with exactly two insertions.
"""
import sys


def transform(text: str) -> str:
    inc_anchor = "#if defined(CUDA)\n#include \"CUDAStream.h\""
    mk_anchor = "#if defined(CUDA)\n  // Use the CUDA implementation"
    if text.count(inc_anchor) != 1 or text.count(mk_anchor) != 1:
        raise SystemExit("StreamModels.h does not look like the BabelStream 5.0 layout; patch it by hand "
                         "(see INTEGRATION.md)")
    text = text.replace(inc_anchor, "#if defined(B200)\n#include \"B200Stream.h\"\n#elif defined(CUDA)\n#include \"CUDAStream.h\"")
    text = text.replace(mk_anchor, "#if defined(B200)\n  // Use the B200 (sm_100a) implementation\n"
                                   "  return std::make_unique<B200Stream<T>>(args...);\n\n"
                                   "#elif defined(CUDA)\n  // Use the CUDA implementation")
    return text


def main():
    src, dst = sys.argv[1], sys.argv[2]
    with open(src) as f:
        out = transform(f.read())
    with open(dst, "w") as f:
        f.write(out)


if __name__ == "__main__":
    main()
