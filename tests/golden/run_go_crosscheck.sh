#!/bin/sh
# Pins the oracle against the real Go standard library: builds oracle/go_crosscheck (stdlib only, no module download)
# wherever a Go toolchain exists, replays the committed vectors and, on success, writes tests/golden/go_<version>.ok.
# The build image of this repository and its GPU boxes have no Go (`go version`: not found), so until someone runs this
# on a machine that has one, parity with the reference stays "unpinned" (DESIGN.md section 2).
set -e
HERE=$(cd "$(dirname "$0")" && pwd)
ROOT=$(cd "$HERE/../.." && pwd)
if ! command -v go >/dev/null 2>&1; then
    echo "run_go_crosscheck: no Go toolchain on this machine (go version: not found); parity stays unpinned" >&2
    exit 3
fi
V=$(go env GOVERSION)
BIN=$(mktemp -d)/go_crosscheck
(cd "$ROOT/oracle/go_crosscheck" && GO111MODULE=off CGO_ENABLED=0 go build -o "$BIN" main.go)
"$BIN" "$HERE/adapter_vectors.json.gz" "$HERE/hand_vectors.json" | tee "$HERE/$V.log"
grep -q '^OK ' "$HERE/$V.log" && mv "$HERE/$V.log" "$HERE/$V.ok" && echo "parity pinned against $V: $HERE/$V.ok"
