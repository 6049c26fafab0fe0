#!/bin/sh
# Regenerates tests/golden/nls_host_reference.txt by running the UNMODIFIED reference (oracle/_ref, built by
# oracle/Makefile from /root/reference) on the reference's own nonlinear test problems. Needs /root/reference.
set -e
cd "$(dirname "$0")/.."
REF=${REF:-/root/reference}
make -C oracle ref >/dev/null
/usr/bin/g++ -std=c++17 -O2 -march=x86-64-v3 -DEIGEN_NO_DEBUG -DFMT_HEADER_ONLY -DNDEBUG -w \
  -I$REF/include -I$REF/dependencies/eigen -I$REF/dependencies/fmt/include -I$REF/dependencies/geometry_utils/include \
  oracle/ref_nls_host_golden.cc -Loracle/_ref -lminiopt_ref -Wl,-rpath,$PWD/oracle/_ref -lpthread -o oracle/_ref/ref_nls_host_golden
oracle/_ref/ref_nls_host_golden > tests/golden/nls_host_reference.txt
wc -l tests/golden/nls_host_reference.txt
