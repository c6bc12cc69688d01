#!/bin/sh
# Regenerates tests/golden/gg/*.{gg,script}: grove streams (format 0.3) written by the UNMODIFIED reference and the
# reference's own answers on the groves it restores from them.  Needs the reference checkout (this container only);
# the generator is ggfast::write_script in tests/cpp/grove_tests.cpp, compiled into ref_parity.bin.
set -e
here=$(cd "$(dirname "$0")" && pwd)
make -C "$here/../cpp" ref
mkdir -p "$here/gg"
"$here/../cpp/ref_parity.bin" --write-gg-fixtures "$here/gg"
