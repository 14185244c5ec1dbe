// Stand-in: the visualiser is not compiled into oracle/_ref.  This header also declares the one hook the build needs:
// FisherInfoManager.cpp reads and writes its lookup table at a hard-coded path under /root/dev_ws; oracle/Makefile replaces that
// string literal (and nothing else) by a call of this function, which the harness points at a temporary file.
#pragma once
#include <string>
std::string fitslam_ref_lookup_path();
