#!/bin/bash
# usage: scripts/fuzz_reader_v3_asan.sh index.idx auxiliary.idx   (host only; needs g++ with libasan)
set -eu
here=$(cd "$(dirname "$0")" && pwd)
g++ -std=c++17 -O1 -g -fsanitize=address,undefined -fno-omit-frame-pointer -o /tmp/lb200_fuzz_reader_v3 "$here/fuzz_reader_v3_asan.cpp"
/tmp/lb200_fuzz_reader_v3 "$1" "$2"
