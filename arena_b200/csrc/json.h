// Minimal JSON reader for the loader: a DOM whose nodes keep the RAW token text
// of scalars (Quantity.UnmarshalJSON works on raw bytes, quotes included:
// vendor/k8s.io/apimachinery/pkg/api/resource/quantity.go:648-667) and decode
// string escapes on demand.  Object keys are matched case-sensitively and the
// LAST duplicate wins, as with the reference's json-iterator configuration
// (vendor/k8s.io/apimachinery/pkg/runtime/serializer/json/json.go:104-113).
#pragma once
#include <cstddef>
#include <string>
#include <string_view>
#include <utility>
#include <vector>

namespace arena_json {

enum class Type { Null, Bool, Number, String, Array, Object };

struct Value {
    Type type = Type::Null;
    std::string_view raw;                                  // token text (scalars: whole token incl. quotes)
    std::vector<Value> items;                              // Array
    std::vector<std::pair<std::string, Value>> members;    // Object, in document order

    bool is_null() const { return type == Type::Null; }
    bool is_object() const { return type == Type::Object; }
    bool is_array() const { return type == Type::Array; }
    bool is_string() const { return type == Type::String; }
    // last member with that key, or nullptr
    const Value *get(std::string_view key) const;
    // decoded string (escapes resolved); empty for non-strings
    std::string str() const;
};

struct ParseError {
    size_t offset = 0;
    std::string message;
};

// Parses exactly one JSON document (surrounding whitespace allowed).
bool parse(std::string_view text, Value &out, ParseError &err);

}  // namespace arena_json
