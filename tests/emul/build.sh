#!/bin/sh
# builds tests/emul/libemul.so (test infrastructure: CPU execution of the kernel phase functions)
set -e
cd "$(dirname "$0")"
g++ -O2 -fPIC -shared -std=c++17 -I../../q-e_b200/csrc -o libemul.so emul_vloc.cpp
