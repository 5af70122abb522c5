// Overlay for the ATProfiler header: its macros exist only under _WIN32 (atp.h:133-293) but
// renderer.cpp:293,308,448-450 uses them. Keep the real types, make the macros no-ops.
#pragma once
#include_next "utilities/ATP/atp.h"
#ifndef _WIN32
#define ATP_REGISTER(name)
#define ATP_REGISTER_M(name, n) static ATP::TestType atp_oracle_slot_##name = {}
#define ATP_GET_TESTTYPE(name) (&atp_oracle_slot_##name)
#define ATP_START(name)
#define ATP_END(name)
#define ATP_BLOCK(name)
#define ATP_BLOCK_M(name, idx)
#endif
