/*
 * Stand-in for 3rdparty/library_suffix/build.h (FWGS/library-suffix @ e9fc1430),
 * a submodule that is absent from the mounted reference tree (.SUBMODULES.json).
 * The reference only needs the platform switches below on this box
 * (x86-64 Linux, little endian); every other XASH_* test defaults to 0.
 * Needed by everything that reads the engine's SDK headers (port.h includes "build.h"): the plugin's host side
 * (rc_host.c), the headless engine stand-in (harness/) and the oracle build of the unmodified reference ref_soft.
 * Inside the engine's own tree the real 3rdparty/library_suffix/build.h takes its place.
 */
#pragma once
#ifndef BUILD_H
#define BUILD_H
#define XASH_LINUX 1
#define XASH_POSIX 1
#define XASH_64BIT 1
#define XASH_AMD64 1
#define XASH_LITTLE_ENDIAN 1
#define XASH_BIG_ENDIAN 0
#endif
