// ingest.h — batch JSON ingest / analysis emit (ingest.cc).  Host only, internal to the library.
#pragma once
#include <cstdint>
#include <string>

#include "../../include/triage.h"

namespace triage {

// userPriorities[msg.UserID] (preprocessor.go:83): 0 = no entry
typedef int (*UserLookup)(const void *ctx, const char *user_id, size_t len);

struct IngestOut {
    uint8_t *arena; uint64_t cap;      // padded arena (16-byte aligned starts, pad fill), nullptr: only sizes
    uint64_t *offsets;                 // n + 1
    uint32_t *lengths;                 // n
    triage_meta *meta;                 // n
    uint8_t *status;                   // n: 0 decoded, 1 rejected (the handler answers 400: bad JSON or a field of the wrong type)
};

// returns "" or an argument error; *total = padded arena bytes needed / written
std::string ingest_json(const uint8_t *bodies, const uint64_t *body_offsets, uint32_t n, int pad, UserLookup lookup, const void *ctx,
                        const IngestOut &out, uint64_t *total);

// {"is_question":b,"sentiment":"s","word_count":n} per result; offsets n + 1; returns bytes needed (nothing written beyond cap)
uint64_t emit_analysis_json(const triage_result *res, uint32_t n, char *out, uint64_t cap, uint64_t *offsets);

}  // namespace triage
