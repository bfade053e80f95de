// Minimal JSON reader for the loader: a DOM whose nodes keep the RAW token text
// of scalars (Quantity.UnmarshalJSON works on raw bytes, quotes included:
// vendor/k8s.io/apimachinery/pkg/api/resource/quantity.go:648-667) and decode
// string escapes on demand.  Object keys are matched case-sensitively and the
// LAST duplicate wins, as with the reference's json-iterator configuration
// (vendor/k8s.io/apimachinery/pkg/runtime/serializer/json/json.go:104-113).
//
// The whole document is always VALIDATED, but only members whose key a caller-
// supplied filter accepts are materialised: a real PodList is ~3 KB per pod of
// which the accounting path reads a few dozen bytes.  Keys are views into the
// input (or into a side arena when they contain escapes), so building the tree
// touches the allocator only when the node pool grows.
#pragma once
#include <cstddef>
#include <cstdint>
#include <deque>
#include <string>
#include <string_view>
#include <utility>
#include <vector>

namespace arena_json {

enum class Type { Null, Bool, Number, String, Array, Object };

// One node of a parsed document.  All nodes of a tree live in ONE contiguous pool
// (Tree::nodes) and link to each other by self-relative indices, so the pool can grow
// (memcpy) while it is being built and a Tree can be reused item after item without
// touching the allocator.
struct Value {
    Type type = Type::Null;
    uint32_t n = 0;                 // number of (kept) children
    int32_t first = 0;              // first child, as an index relative to this node (0 = none)
    int32_t next = 0;               // next sibling, relative to this node (0 = none)
    std::string_view raw;           // token text (scalars: whole token incl. quotes; containers: whole span)
    std::string_view key;           // member key when the parent is an object (decoded)

    bool is_null() const { return type == Type::Null; }
    bool is_object() const { return type == Type::Object; }
    bool is_array() const { return type == Type::Array; }
    bool is_string() const { return type == Type::String; }
    size_t size() const { return n; }
    const Value *first_child() const { return first ? this + first : nullptr; }
    const Value *next_sibling() const { return next ? this + next : nullptr; }

    struct Iter {
        const Value *v;
        const Value &operator*() const { return *v; }
        Iter &operator++() { v = v->next_sibling(); return *this; }
        bool operator!=(const Iter &o) const { return v != o.v; }
    };
    struct Children {
        const Value *head;
        Iter begin() const { return {head}; }
        Iter end() const { return {nullptr}; }
    };
    // array elements / object members (kept ones), in document order
    Children children() const { return {first_child()}; }
    // last member with that key, or nullptr
    const Value *get(std::string_view key) const;
    // decoded string (escapes resolved); empty for non-strings
    std::string str() const;
};

struct Tree {
    std::vector<Value> nodes;                   // nodes[0] is the root
    std::deque<std::string> keys;               // decoded keys that contained escapes
    const Value &root() const { return nodes.front(); }
    void clear() { nodes.clear(); keys.clear(); }
};

struct ParseError {
    size_t offset = 0;
    std::string message;
};

// Decides which object members are materialised.  `parent_key` is the key under
// which the enclosing object sits ("" for the document root; elements of an array
// inherit the array's key).  Returning false validates and drops the member.
using KeyFilter = bool (*)(std::string_view parent_key, std::string_view key);

struct Range {
    size_t begin = 0, end = 0;                  // [begin, end) byte offsets into the document
};

struct ParseOptions {
    KeyFilter filter = nullptr;                 // nullptr keeps everything
    // List envelopes: when the ROOT object has a member `split_key` whose value is an
    // array, do not descend into it: record the byte range of every element in
    // `split_out` (to be parsed, possibly in parallel, with parse_range) and leave the
    // member as an Array with no items.  If the array cannot be split by the structural
    // scan (scalar elements, damaged text, a duplicate `split_key`), it is parsed in
    // place as usual and `*split_done` stays false.
    std::string_view split_key;
    std::vector<Range> *split_out = nullptr;
    bool *split_done = nullptr;
    // Someone else (the GPU structural scan) already knows where containers end: when set,
    // this is asked for the end of the array opening at `open` instead of scanning for it,
    // and `split_out` receives ONE range, the array itself; its elements are the caller's
    // business (it must validate them).
    bool (*split_hook)(void *user, size_t open, size_t *end) = nullptr;
    void *split_user = nullptr;
};

// Parses exactly one JSON document (surrounding whitespace allowed) into `out`
// (cleared first).  The tree refers to `text`, which must outlive it.
bool parse(std::string_view text, Tree &out, ParseError &err, const ParseOptions &opt = ParseOptions());

// Parses the one value that fills text[r.begin, r.end) (whitespace around it allowed)
// as if it sat `depth` levels inside the document under the key `parent_key`.
// Error offsets are relative to `text`.
bool parse_range(std::string_view text, Range r, int depth, std::string_view parent_key, Tree &out, ParseError &err,
                 const ParseOptions &opt);

// Structural scan (64 bytes per step: quote / backslash / bracket bitmasks): element
// ranges of the array whose '[' is text[open].  Returns false when the array is not
// cleanly splittable; a true result promises only that, if every range parses as one
// value, text[open, end) is a valid array.
bool split_array(std::string_view text, size_t open, std::vector<Range> &out, size_t &end);

}  // namespace arena_json
