// TEST INFRASTRUCTURE ONLY (oracle build).  The fork's host library declares a
// 10-argument BlackBoxJam (foldedmv-offload.h:221) while every non-TMR top.cpp
// defines the 9-argument one (cnvW1A1/hw/top.cpp:238-240): SURVEY.md App. C.2.
// This forwarder lets the unmodified sources link.
#include "ap_int.h"
void BlackBoxJam(ap_uint<64> *in, ap_uint<64> *out, unsigned int doInit, unsigned int targetLayer, unsigned int targetMem,
                 unsigned int targetInd, unsigned int targetThresh, ap_uint<64> val, unsigned int numReps);
void BlackBoxJam(ap_uint<64> *in, ap_uint<64> *out, unsigned int doInit, unsigned int targetLayer, unsigned int targetMem,
                 unsigned int targetInd, unsigned int targetThresh, ap_uint<64> val, unsigned int numReps, unsigned int) {
  BlackBoxJam(in, out, doInit, targetLayer, targetMem, targetInd, targetThresh, val, numReps);
}
