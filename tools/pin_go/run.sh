#!/bin/bash
# Pin the oracle against Go: needs a Go toolchain (>= 1.23) and module access (or a populated module cache).
#   tools/pin_go/run.sh        -> tests/golden/go/*.txt, *.hdr ; then: python -m pytest tests/test_go_pinned.py
set -euo pipefail
here="$(cd "$(dirname "$0")" && pwd)"
root="$(cd "$here/../.." && pwd)"
python "$root/tools/pin_go/make_inputs.py"          # (re)writes tests/golden/go_inputs.json — deterministic
cd "$here"
go mod tidy                                          # resolves the indirect requirements at the pinned versions
go run . "$root"
ls -l "$root/tests/golden/go"
