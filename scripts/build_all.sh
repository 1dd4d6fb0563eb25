#!/bin/bash
# builds both libraries and the C++ test programs in-tree; stops at the first failure
set -e
R=$(cd "$(dirname "$0")/.." && pwd)
make -s -j8 -C "$R/titan_b200/csrc" 2>&1 | grep -E "error|warning: v" || true
make -s -C "$R/titan_b200/csrc" >/dev/null
make -s -C "$R/titan_b200/host"
make -s -C "$R/tests/cpp"
echo "built: $(ls -la --time-style=+%T "$R/titan_b200/libtitan_b200.so" | awk '{print $6}')"
