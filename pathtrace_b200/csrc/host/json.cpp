#include "json.h"

#include <cstdint>
#include <cstdlib>
#include <cstring>

namespace b200pt {

static const Json g_null;
static const std::vector<Json> g_no_items;

const Json *Json::find(const char *key) const
{
    if (type_ != Object)
        return nullptr;
    const Json *found = nullptr;
    for (const auto &m : *object_)
        if (m.first == key)
            found = &m.second; // a repeated key: the last one counts, as in nlohmann
    return found;
}

const Json &Json::operator[](const char *key) const
{
    const Json *j = find(key);
    return j ? *j : g_null;
}

const Json &Json::at(size_t i) const
{
    if (type_ != Array || i >= array_->size())
        throw std::runtime_error("json: array index out of range");
    return (*array_)[i];
}

const std::vector<Json> &Json::items() const { return type_ == Array ? *array_ : g_no_items; }

double Json::as_double() const
{
    if (type_ == Number)
        return number_;
    if (type_ == Bool)
        return bool_ ? 1.0 : 0.0;
    throw std::runtime_error("json: value is not a number");
}

bool Json::as_bool() const
{
    if (type_ == Bool)
        return bool_;
    if (type_ == Number)
        return number_ != 0.0;
    throw std::runtime_error("json: value is not a boolean");
}

const std::string &Json::as_string() const
{
    if (type_ != String)
        throw std::runtime_error("json: value is not a string");
    return *string_;
}

class JsonParser {
public:
    explicit JsonParser(const std::string &s) : s_(s), i_(0) {}

    Json parse_document()
    {
        Json v = parse_value();
        skip_ws();
        if (i_ != s_.size())
            fail("trailing characters");
        return v;
    }

private:
    const std::string &s_;
    size_t i_;

    [[noreturn]] void fail(const char *what)
    {
        throw std::runtime_error(std::string("json: ") + what + " at byte " + std::to_string(i_));
    }

    void skip_ws()
    {
        while (i_ < s_.size() && (s_[i_] == ' ' || s_[i_] == '\t' || s_[i_] == '\n' || s_[i_] == '\r'))
            i_++;
    }

    Json parse_value()
    {
        skip_ws();
        if (i_ >= s_.size())
            fail("unexpected end");
        char c = s_[i_];
        if (c == '{')
            return parse_object();
        if (c == '[')
            return parse_array();
        if (c == '"') {
            Json v;
            v.type_ = Json::String;
            v.string_.reset(new std::string(parse_string()));
            return v;
        }
        if (c == 't' || c == 'f' || c == 'n')
            return parse_literal();
        return parse_number();
    }

    Json parse_literal()
    {
        Json v;
        if (s_.compare(i_, 4, "true") == 0) {
            v.type_ = Json::Bool;
            v.bool_ = true;
            i_ += 4;
        } else if (s_.compare(i_, 5, "false") == 0) {
            v.type_ = Json::Bool;
            v.bool_ = false;
            i_ += 5;
        } else if (s_.compare(i_, 4, "null") == 0) {
            i_ += 4;
        } else {
            fail("bad literal");
        }
        return v;
    }

    Json parse_number()
    {
        const char *begin = s_.c_str() + i_;
        // Plain decimals (every number of a scene file) without strtod: with at most 15 significant digits the
        // mantissa is exact in a double, 10^|e| is exact up to 22, so one multiplication or division gives the
        // correctly rounded value -- the same bits strtod returns.  Anything else goes to strtod.
        {
            const char *p = begin;
            const bool neg = *p == '-';
            if (neg)
                p++;
            uint64_t mant = 0;
            int digits = 0, exp10 = 0;
            const char *first = p;
            while (*p >= '0' && *p <= '9') {
                mant = mant * 10 + (uint64_t)(*p - '0');
                digits += (mant != 0);
                p++;
            }
            bool ok = p != first;
            if (*p == '.') {
                p++;
                const char *frac = p;
                while (*p >= '0' && *p <= '9') {
                    mant = mant * 10 + (uint64_t)(*p - '0');
                    digits += (mant != 0);
                    exp10--;
                    p++;
                }
                ok = ok && p != frac;
            }
            if (ok && (*p == 'e' || *p == 'E')) {
                const char *q = p + 1;
                const bool eneg = *q == '-';
                if (*q == '-' || *q == '+')
                    q++;
                int e = 0;
                const char *efirst = q;
                while (*q >= '0' && *q <= '9' && e < 1000) {
                    e = e * 10 + (*q - '0');
                    q++;
                }
                if (q == efirst || (*q >= '0' && *q <= '9'))
                    ok = false;
                exp10 += eneg ? -e : e;
                p = q;
            }
            if (ok && digits <= 15 && exp10 >= -22 && exp10 <= 22) {
                static const double pow10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                                 1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
                double d = (double)mant;
                d = exp10 < 0 ? d / pow10[-exp10] : d * pow10[exp10];
                i_ += (size_t)(p - begin);
                Json v;
                v.type_ = Json::Number;
                v.number_ = neg ? -d : d;
                return v;
            }
        }
        char *end = nullptr;
        double d = strtod(begin, &end);
        if (end == begin)
            fail("bad number");
        i_ += (size_t)(end - begin);
        Json v;
        v.type_ = Json::Number;
        v.number_ = d;
        return v;
    }

    std::string parse_string()
    {
        std::string out;
        i_++; // opening quote
        while (true) {
            if (i_ >= s_.size())
                fail("unterminated string");
            char c = s_[i_++];
            if (c == '"')
                break;
            if (c != '\\') {
                out.push_back(c);
                continue;
            }
            if (i_ >= s_.size())
                fail("bad escape");
            char e = s_[i_++];
            switch (e) {
            case 'n': out.push_back('\n'); break;
            case 't': out.push_back('\t'); break;
            case 'r': out.push_back('\r'); break;
            case 'b': out.push_back('\b'); break;
            case 'f': out.push_back('\f'); break;
            case 'u': {
                if (i_ + 4 > s_.size())
                    fail("bad \\u escape");
                unsigned cp = (unsigned)strtoul(s_.substr(i_, 4).c_str(), nullptr, 16);
                i_ += 4;
                if (cp < 0x80) {
                    out.push_back((char)cp);
                } else if (cp < 0x800) {
                    out.push_back((char)(0xC0 | (cp >> 6)));
                    out.push_back((char)(0x80 | (cp & 0x3F)));
                } else {
                    out.push_back((char)(0xE0 | (cp >> 12)));
                    out.push_back((char)(0x80 | ((cp >> 6) & 0x3F)));
                    out.push_back((char)(0x80 | (cp & 0x3F)));
                }
                break;
            }
            default: out.push_back(e); break; // \" \\ \/
            }
        }
        return out;
    }

    Json parse_array()
    {
        Json v;
        v.type_ = Json::Array;
        v.array_.reset(new std::vector<Json>());
        v.array_->reserve(4); // most arrays are vec3s: one allocation instead of three
        i_++;
        skip_ws();
        if (i_ < s_.size() && s_[i_] == ']') {
            i_++;
            return v;
        }
        while (true) {
            v.array_->push_back(parse_value());
            skip_ws();
            if (i_ >= s_.size())
                fail("unterminated array");
            if (s_[i_] == ',') {
                i_++;
                continue;
            }
            if (s_[i_] == ']') {
                i_++;
                break;
            }
            fail("expected , or ]");
        }
        return v;
    }

    Json parse_object()
    {
        Json v;
        v.type_ = Json::Object;
        v.object_.reset(new Json::Members());
        v.object_->reserve(4);
        i_++;
        skip_ws();
        if (i_ < s_.size() && s_[i_] == '}') {
            i_++;
            return v;
        }
        while (true) {
            skip_ws();
            if (i_ >= s_.size() || s_[i_] != '"')
                fail("expected key");
            std::string key = parse_string();
            skip_ws();
            if (i_ >= s_.size() || s_[i_] != ':')
                fail("expected :");
            i_++;
            v.object_->emplace_back(std::move(key), parse_value());
            skip_ws();
            if (i_ >= s_.size())
                fail("unterminated object");
            if (s_[i_] == ',') {
                i_++;
                continue;
            }
            if (s_[i_] == '}') {
                i_++;
                break;
            }
            fail("expected , or }");
        }
        return v;
    }
};

Json Json::parse(const std::string &text)
{
    JsonParser p(text);
    return p.parse_document();
}

} // namespace b200pt
