// regex_vm.h — AddKeywordPattern for patterns that are not plain literals (internal/preprocessor/preprocessor.go:176-184
// takes any regexp.Compile-able string): a subset of RE2 syntax compiled to a small rune-level program that the device
// runs as an ordered-thread (Pike) simulation, one message per thread, counting what
// len(pattern.FindAllString(content, -1)) counts (:146): successive leftmost-first, non-overlapping matches.
//
// Subset: optional leading flag group "(?i)", "(?s)" or both; literals (UTF-8); escapes of metacharacters, \n \t \r \f \v,
// \xHH, \d \D \w \W \s \S (ASCII, as in RE2); "." (any rune but \n unless (?s)); classes [..] with ranges, negation and
// the escapes above; groups ( ) and (?: ); alternation; ? * + {n} {n,} {n,m} and their lazy forms.  Outside the subset
// (an error, never a silent approximation): anchors and \b, named groups, POSIX classes, flags inside the pattern,
// non-ASCII letters under (?i), U+FFFD, patterns that can match the empty string, a repeated part that can be empty,
// more than kRegexMaxInsts instructions after expansion.
//
// Everything is defined on Go's runes: invalid UTF-8 bytes decode to U+FFFD of width 1, which "." and negated classes
// match and literals never do.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace triage {

constexpr uint32_t kRegexMaxInsts = 64;    // thread sets are 64-bit masks on the device
constexpr uint32_t kRegexMaxRepeat = 100;

enum RegexOp : uint8_t { ROP_RUNE = 0, ROP_SPLIT = 1, ROP_JMP = 2, ROP_MATCH = 3 };

struct RegexInst {      // 8 bytes
    uint8_t op;
    uint8_t pad;
    uint16_t cls;       // ROP_RUNE: class index (global, over all programs of the dictionary)
    uint16_t x, y;      // ROP_SPLIT: x preferred, then y; ROP_JMP: x  (indices inside the program)
};

struct RegexClass {     // a set of runes: 128-bit ASCII bitmap + sorted ranges for the rest, optionally negated
    uint32_t ascii[4];
    uint32_t first_range, n_ranges;   // into the range table, (lo, hi) pairs, runes >= 0x80
    uint32_t negate;
    uint32_t pad;
};

struct RegexProgram {
    uint32_t first_inst, n_inst;
    uint32_t prio;      // 1..4
    uint32_t pad;
};

// host side: all regex patterns of a dictionary
struct RegexSet {
    std::vector<RegexProgram> programs;
    std::vector<RegexInst> insts;
    std::vector<RegexClass> classes;
    std::vector<uint32_t> ranges;     // lo, hi, lo, hi, ...
    std::vector<std::string> sources;
    void clear() { programs.clear(); insts.clear(); classes.clear(); ranges.clear(); sources.clear(); }
    bool empty() const { return programs.empty(); }
};

// Compiles `pattern` and appends it to `set`.  Returns "" or the reason it is outside the subset.
std::string regex_compile(const char *pattern, uint32_t len, uint32_t prio, RegexSet *set);

// Host reference of the device walk (same program, same thread order): used by the host checks in tests/csrc.
uint32_t regex_count_host(const RegexSet &set, uint32_t program, const uint8_t *s, uint64_t n);

}  // namespace triage
