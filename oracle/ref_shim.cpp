// TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// extern "C" forwarders around the UNMODIFIED reference translation unit
// /root/reference/utils/pafprocess/pafprocess.cpp, which exports C++-mangled symbols and is
// normally reached through SWIG (pafprocess.i:14).  SWIG is not in this image, so the checker
// binds these seven plain-C names with ctypes instead.  Nothing is copied from the reference:
// the build recipe (oracle/build_oracle.py) compiles pafprocess.cpp where it lies and links
// this file next to it into oracle/_ref/libpafprocess_ref.so (git-ignored).
#include "pafprocess.h"  // resolved with -I/root/reference/utils/pafprocess at build time

extern "C" {
int ref_process_paf(int p1, int p2, int p3, float *peaks, int f1, int f2, int f3, float *pafmap,
                    int min_img_size) {
    return process_paf(p1, p2, p3, peaks, f1, f2, f3, pafmap, min_img_size);  // pafprocess.h:70
}
int ref_get_num_humans() { return get_num_humans(); }                                  // :71
int ref_get_part_peak_id(int human, int part) { return get_part_peak_id(human, part); }  // :72
float ref_get_score(int human) { return get_score(human); }                            // :73
int ref_get_part_x(int cid) { return get_part_x(cid); }                                // :74
int ref_get_part_y(int cid) { return get_part_y(cid); }                                // :75
float ref_get_part_score(int cid) { return get_part_score(cid); }                      // :76
}
