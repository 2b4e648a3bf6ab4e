#pragma once
// A small strict-JSON DOM reader written for the scene loader.
//
// The reference parses scenes with rapidjson 1.1.0 (vendor/rapidjson @24b5e7a, default parse flags) and reads
// every number through GenericValue::GetFloat() == float(GetDouble()).  rapidjson's default ("normal precision")
// number conversion is NOT a correctly rounded strtod: it accumulates the significand (uint64, then double) and
// applies one multiply or divide by a power of ten (reader.h ParseNumber, internal/strtod.h
// StrtodNormalPrecision/FastPath).  To hand the kernels bit-identical inputs we restate that published
// algorithm here (number grammar, int/double classification, significand accumulation, 10^p scaling) instead of
// calling strtod.  Everything else is ordinary RFC 8259 parsing with the predicates the loader needs
// (is_int == rapidjson IsInt, first-match member lookup == FindMember).
//
// Numeric arrays (vertices / triangles / uvs: millions of entries) are stored unboxed.

#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <string>
#include <string_view>
#include <utility>
#include <vector>

namespace crt::jsonreader {

struct Number {
    double value = 0.0;
    bool is_int = false; // integer lexeme within int32 range (rapidjson kIntFlag)
};

struct Value {
    enum class Kind : uint8_t { Null, Bool, Number, String, Array, NumberArray, Object };
    Kind kind = Kind::Null;
    bool boolean = false;
    Number number;
    std::string string;
    std::vector<Value> items;                             // Kind::Array
    std::vector<double> numbers;                          // Kind::NumberArray
    bool numbers_all_int = true;                          // every element of `numbers` is_int
    std::vector<std::pair<std::string, Value>> members;   // Kind::Object, file order

    bool is_object() const { return kind == Kind::Object; }
    bool is_array() const { return kind == Kind::Array || kind == Kind::NumberArray; }
    bool is_number() const { return kind == Kind::Number; }
    bool is_int() const { return kind == Kind::Number && number.is_int; }
    bool is_bool() const { return kind == Kind::Bool; }
    bool is_string() const { return kind == Kind::String; }
    size_t size() const { return kind == Kind::NumberArray ? numbers.size() : items.size(); }
    bool empty() const { return size() == 0; }
    float as_float() const { return static_cast<float>(number.value); }
    int as_int() const { return static_cast<int>(number.value); }

    // first member with that name, like rapidjson FindMember
    const Value *find(std::string_view name) const {
        if (kind != Kind::Object)
            return nullptr;
        for (const auto &m : members)
            if (m.first == name)
                return &m.second;
        return nullptr;
    }
};

namespace detail {

inline const double *pow10_table() {
    static double table[309];
    static bool ready = false;
    if (!ready) {
        char buf[16];
        for (int i = 0; i <= 308; ++i) {
            std::snprintf(buf, sizeof buf, "1e%d", i);
            table[i] = std::strtod(buf, nullptr); // correctly rounded, same as the literals 1e+0 .. 1e+308
        }
        ready = true;
    }
    return table;
}

inline double scale_pow10(double significand, int exp) { // FastPath
    if (exp < -308)
        return 0.0;
    if (exp >= 0)
        return significand * pow10_table()[exp];
    return significand / pow10_table()[-exp];
}

class Parser {
public:
    Parser(const char *begin, const char *end) : p_(begin), end_(end) {}

    bool parse_document(Value &out) {
        skip_ws();
        if (p_ == end_)
            return false; // empty document
        if (!parse_value(out, 0))
            return false;
        skip_ws();
        return p_ == end_; // the root must not be followed by other values
    }

private:
    const char *p_, *end_;
    static constexpr int kMaxDepth = 512;

    int peek() const { return p_ < end_ ? static_cast<unsigned char>(*p_) : -1; }
    static bool digit(int c) { return c >= '0' && c <= '9'; }
    void skip_ws() {
        while (p_ < end_ && (*p_ == ' ' || *p_ == '\n' || *p_ == '\r' || *p_ == '\t'))
            ++p_;
    }
    bool literal(const char *word) {
        size_t n = std::strlen(word);
        if (size_t(end_ - p_) < n || std::memcmp(p_, word, n) != 0)
            return false;
        p_ += n;
        return true;
    }

    bool parse_value(Value &out, int depth) {
        if (depth > kMaxDepth)
            return false;
        switch (peek()) {
            case 'n': out.kind = Value::Kind::Null; return literal("null");
            case 't': out.kind = Value::Kind::Bool; out.boolean = true; return literal("true");
            case 'f': out.kind = Value::Kind::Bool; out.boolean = false; return literal("false");
            case '"': out.kind = Value::Kind::String; return parse_string(out.string);
            case '[': return parse_array(out, depth);
            case '{': return parse_object(out, depth);
            default: out.kind = Value::Kind::Number; return parse_number(out.number);
        }
    }

    bool parse_number(Number &out) {
        const bool minus = peek() == '-';
        if (minus)
            ++p_;
        uint64_t i64 = 0;
        double d = 0.0;
        bool use_double = false;
        int significand_digits = 0;
        if (peek() == '0') {
            ++p_;
        } else if (peek() >= '1' && peek() <= '9') {
            i64 = uint64_t(*p_++ - '0');
            const uint64_t limit = minus ? 0x0CCCCCCCCCCCCCCCull : 0x1999999999999999ull;
            const int last = minus ? '8' : '5';
            while (digit(peek())) {
                if (i64 >= limit && (i64 != limit || peek() > last)) {
                    d = static_cast<double>(i64);
                    use_double = true;
                    break;
                }
                i64 = i64 * 10 + uint64_t(*p_++ - '0');
                ++significand_digits;
            }
        } else {
            return false;
        }
        if (use_double)
            while (digit(peek()))
                d = d * 10 + (*p_++ - '0');

        bool integer_lexeme = !use_double;
        int exp_frac = 0;
        if (peek() == '.') {
            ++p_;
            if (!digit(peek()))
                return false;
            integer_lexeme = false;
            if (!use_double) {
                while (digit(peek())) {
                    if (i64 > 0x1FFFFFFFFFFFFFull) // 2^53 - 1
                        break;
                    i64 = i64 * 10 + uint64_t(*p_++ - '0');
                    --exp_frac;
                    if (i64 != 0)
                        ++significand_digits;
                }
                d = static_cast<double>(i64);
                use_double = true;
            }
            while (digit(peek())) {
                if (significand_digits < 17) {
                    d = d * 10.0 + (*p_++ - '0');
                    --exp_frac;
                    if (d > 0.0)
                        ++significand_digits;
                } else {
                    ++p_;
                }
            }
        }
        int exp = 0;
        if (peek() == 'e' || peek() == 'E') {
            ++p_;
            integer_lexeme = false;
            if (!use_double) {
                d = static_cast<double>(i64);
                use_double = true;
            }
            bool exp_minus = false;
            if (peek() == '+')
                ++p_;
            else if (peek() == '-') {
                exp_minus = true;
                ++p_;
            }
            if (!digit(peek()))
                return false;
            exp = *p_++ - '0';
            if (exp_minus) {
                const int max_exp = (exp_frac + 2147483639) / 10;
                while (digit(peek())) {
                    exp = exp * 10 + (*p_++ - '0');
                    if (exp > max_exp)
                        while (digit(peek()))
                            ++p_;
                }
                exp = -exp;
            } else {
                const int max_exp = 308 - exp_frac;
                while (digit(peek())) {
                    exp = exp * 10 + (*p_++ - '0');
                    if (exp > max_exp)
                        return false; // number too big
                }
            }
        }
        if (use_double) {
            const int p = exp + exp_frac;
            if (p < -308) {
                d = scale_pow10(d, -308);
                d = scale_pow10(d, p + 308);
            } else {
                d = scale_pow10(d, p);
            }
            if (d > 1.7976931348623157e308)
                return false;
            out.value = minus ? -d : d;
            out.is_int = false;
        } else {
            // integer lexeme: "-0" is the int 0 (positive zero once converted), exactly as rapidjson's Int(0)
            out.value = minus ? -static_cast<double>(i64) : static_cast<double>(i64);
            if (minus && i64 == 0)
                out.value = 0.0;
            out.is_int = integer_lexeme && (minus ? i64 <= 2147483648ull : i64 <= 2147483647ull);
        }
        return true;
    }

    static void append_utf8(std::string &s, uint32_t cp) {
        if (cp < 0x80) {
            s.push_back(char(cp));
        } else if (cp < 0x800) {
            s.push_back(char(0xC0 | (cp >> 6)));
            s.push_back(char(0x80 | (cp & 0x3F)));
        } else if (cp < 0x10000) {
            s.push_back(char(0xE0 | (cp >> 12)));
            s.push_back(char(0x80 | ((cp >> 6) & 0x3F)));
            s.push_back(char(0x80 | (cp & 0x3F)));
        } else {
            s.push_back(char(0xF0 | (cp >> 18)));
            s.push_back(char(0x80 | ((cp >> 12) & 0x3F)));
            s.push_back(char(0x80 | ((cp >> 6) & 0x3F)));
            s.push_back(char(0x80 | (cp & 0x3F)));
        }
    }

    bool hex4(uint32_t &out) {
        if (end_ - p_ < 4)
            return false;
        out = 0;
        for (int k = 0; k < 4; ++k) {
            int c = static_cast<unsigned char>(*p_++);
            out <<= 4;
            if (c >= '0' && c <= '9') out |= uint32_t(c - '0');
            else if (c >= 'a' && c <= 'f') out |= uint32_t(c - 'a' + 10);
            else if (c >= 'A' && c <= 'F') out |= uint32_t(c - 'A' + 10);
            else return false;
        }
        return true;
    }

    bool parse_string(std::string &out) {
        ++p_; // opening quote
        out.clear();
        for (;;) {
            if (p_ >= end_)
                return false;
            unsigned char c = static_cast<unsigned char>(*p_++);
            if (c == '"')
                return true;
            if (c < 0x20)
                return false;
            if (c != '\\') {
                out.push_back(char(c));
                continue;
            }
            if (p_ >= end_)
                return false;
            switch (*p_++) {
                case '"': out.push_back('"'); break;
                case '\\': out.push_back('\\'); break;
                case '/': out.push_back('/'); break;
                case 'b': out.push_back('\b'); break;
                case 'f': out.push_back('\f'); break;
                case 'n': out.push_back('\n'); break;
                case 'r': out.push_back('\r'); break;
                case 't': out.push_back('\t'); break;
                case 'u': {
                    uint32_t cp;
                    if (!hex4(cp))
                        return false;
                    if (cp >= 0xD800 && cp <= 0xDBFF) { // high surrogate: a low one must follow
                        uint32_t lo;
                        if (end_ - p_ < 2 || p_[0] != '\\' || p_[1] != 'u')
                            return false;
                        p_ += 2;
                        if (!hex4(lo) || lo < 0xDC00 || lo > 0xDFFF)
                            return false;
                        cp = 0x10000 + (((cp - 0xD800) << 10) | (lo - 0xDC00));
                    } else if (cp >= 0xDC00 && cp <= 0xDFFF) {
                        return false;
                    }
                    append_utf8(out, cp);
                    break;
                }
                default: return false;
            }
        }
    }

    bool parse_array(Value &out, int depth) {
        ++p_; // '['
        out.kind = Value::Kind::NumberArray; // optimistic: demoted to Array on the first non-number
        out.numbers_all_int = true;
        skip_ws();
        if (peek() == ']') {
            ++p_;
            return true;
        }
        for (;;) {
            skip_ws();
            const int c = peek();
            const bool numeric = c == '-' || digit(c);
            if (out.kind == Value::Kind::NumberArray && numeric) {
                Number n;
                if (!parse_number(n))
                    return false;
                out.numbers.push_back(n.value);
                out.numbers_all_int = out.numbers_all_int && n.is_int;
            } else {
                if (out.kind == Value::Kind::NumberArray)
                    demote(out);
                out.items.emplace_back();
                if (!parse_value(out.items.back(), depth + 1))
                    return false;
            }
            skip_ws();
            if (peek() == ',') {
                ++p_;
                continue;
            }
            if (peek() == ']') {
                ++p_;
                return true;
            }
            return false;
        }
    }

    // a mixed array: re-box the numbers read so far (rare: only tiny arrays in scene files)
    static void demote(Value &out) {
        out.kind = Value::Kind::Array;
        out.items.reserve(out.numbers.size() + 1);
        for (double v : out.numbers) {
            Value boxed;
            boxed.kind = Value::Kind::Number;
            boxed.number.value = v;
            // per-element is_int is only known when all were ints; no loader rule accepts a mixed array
            // on the strength of is_int (vectors need 3 numbers, index arrays need all ints), so this is enough
            boxed.number.is_int = out.numbers_all_int;
            out.items.push_back(std::move(boxed));
        }
        out.numbers.clear();
    }

    bool parse_object(Value &out, int depth) {
        ++p_; // '{'
        out.kind = Value::Kind::Object;
        skip_ws();
        if (peek() == '}') {
            ++p_;
            return true;
        }
        for (;;) {
            skip_ws();
            if (peek() != '"')
                return false;
            std::string key;
            if (!parse_string(key))
                return false;
            skip_ws();
            if (peek() != ':')
                return false;
            ++p_;
            skip_ws();
            out.members.emplace_back(std::move(key), Value{});
            if (!parse_value(out.members.back().second, depth + 1))
                return false;
            skip_ws();
            if (peek() == ',') {
                ++p_;
                continue;
            }
            if (peek() == '}') {
                ++p_;
                return true;
            }
            return false;
        }
    }
};

} // namespace detail

inline bool parse(const char *text, size_t length, Value &out) {
    // rapidjson's UTF-8 input stream does not skip a byte-order mark unless asked to; neither do we.
    detail::Parser parser{ text, text + length };
    return parser.parse_document(out);
}

} // namespace crt::jsonreader
