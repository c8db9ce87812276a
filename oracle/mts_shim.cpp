/* ORACLE (test infrastructure, not product code).
 *
 * Table-injection stand-in for the MTS-ESP client library.  The reference
 * engine pulls its 128 note frequencies through exactly three symbols
 * (reference src/tuning.cpp:25-34 -> libs/MTS-ESP/Client/libMTSClient.h:125-154);
 * this shim serves them from a caller-set table so that the reference can be
 * driven headless without a libMTS master.
 */
#include "libMTSClient.h"
#include <cmath>

static double g_table[128];
static bool g_table_set = false;

struct MTSClient {
    int dummy;
};
static MTSClient g_client;

extern "C" void ref_mts_set_table(const double *freqs128)
{
    for (int i = 0; i < 128; ++i)
        g_table[i] = freqs128[i];
    g_table_set = true;
}

MTSClient *MTS_RegisterClient() { return &g_client; }
void MTS_DeregisterClient(MTSClient *) {}

double MTS_NoteToFrequency(MTSClient *, char midinote, char)
{
    int n = (int)midinote & 127;
    if (g_table_set)
        return g_table[n];
    /* 12-TET, A4 = 440 Hz: what libMTSClient serves with no master connected. */
    return 440.0 * std::pow(2.0, (n - 69) / 12.0);
}
