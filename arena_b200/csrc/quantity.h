// Host-side resource.Quantity: string -> int64Amount{value, scale} (fast path)
// or a pre-resolved Value() (slow path, inf.Dec in the reference).
// Follows vendor/k8s.io/apimachinery/pkg/api/resource/quantity.go:159-390,
// 648-718, suffix.go:88-198, math.go:59-203, scale_int.go:43-95 and
// vendor/gopkg.in/inf.v0/dec.go:243-335, 455-537.
#pragma once
#include <cstdint>
#include <string_view>

namespace arena_host {

struct ParsedQuantity {
    bool slow = false;     // went through the inf.Dec path
    int64_t amount = 0;    // fast: int64Amount.value ; slow: Value() already resolved
    int32_t scale = 0;     // fast: int64Amount.scale ; slow: 0
    int64_t value = 0;     // Quantity.Value()
};

// ParseQuantity(str).  false = the reference returns an error.
bool parse_quantity(std::string_view s, ParsedQuantity &out);
// Quantity.UnmarshalJSON on the raw token bytes (quotes included for strings).
bool unmarshal_quantity_json(std::string_view raw, ParsedQuantity &out);
// Value() of int64Amount{value, scale} (AsScaledInt64(0) with `ok` dropped).
int64_t amount_value(int64_t value, int32_t scale);
// Column encoding: {amount, scale} with |scale| <= 18 that the kernels decode
// to exactly q.value.
void column_encoding(const ParsedQuantity &q, int64_t &val, int8_t &scale);

}  // namespace arena_host
