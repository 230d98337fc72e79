#!/bin/sh
# builds tests/emul/libemul.so (test infrastructure: CPU execution of the kernel phase functions)
set -e
cd "$(dirname "$0")"
# up to date? (any kernel header newer than the library forces a rebuild)
if [ -f libemul.so ] && [ -z "$(find emul_vloc.cpp build.sh ../../q-e_b200/csrc -newer libemul.so -type f \( -name '*.cuh' -o -name '*.h' -o -name '*.cpp' -o -name '*.sh' \) | head -1)" ]; then exit 0; fi
g++ -O2 -fPIC -shared -std=c++17 -I../../q-e_b200/csrc -o libemul.so emul_vloc.cpp
