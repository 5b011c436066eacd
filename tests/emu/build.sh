#!/bin/sh
# builds the CPU debug harness tests/emu/libref_cuda_emu.so (rc_host.c + rc_emu.cpp). Not a product artefact.
set -e
cd "$(dirname "$0")/../.."
R=${REF:-/root/reference}
DEF="-DHAVE_TGMATH_H=1 -DHAVE_STRCASECMP=1 -DHAVE_STRNLEN=1 -DHAVE_STRCASESTR=1 -DHAVE_STRCHRNUL=1"
INC="-Iinclude/shim -Iinclude -Ixash3d-fwgs_b200/csrc -I$R/engine -I$R/engine/common -I$R/engine/common/imagelib -I$R/filesystem -I$R/public -I$R/common -I$R/pm_shared"
mkdir -p xash3d-fwgs_b200/build
gcc -std=gnu11 -O2 -g -fPIC -Wall -Wno-unused-function -fvisibility=hidden -ffp-contract=off $DEF $INC -c xash3d-fwgs_b200/csrc/rc_host.c -o xash3d-fwgs_b200/build/rc_host.o
g++ -std=c++17 -O2 -g -fPIC -ffp-contract=off -Wall -Wno-unused-function -Ixash3d-fwgs_b200/csrc -c tests/emu/rc_emu.cpp -o xash3d-fwgs_b200/build/rc_emu.o
g++ -shared xash3d-fwgs_b200/build/rc_emu.o xash3d-fwgs_b200/build/rc_host.o xash3d-fwgs_b200/build/public_xash3d_mathlib.o xash3d-fwgs_b200/build/public_matrixlib.o -lm -o tests/emu/libref_cuda_emu.so
