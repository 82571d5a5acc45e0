#include "json.h"

#include <cstdlib>
#include <cstring>

namespace b200pt {

static const Json g_null;

const Json &Json::operator[](const std::string &key) const
{
    if (type_ != Object)
        return g_null;
    auto it = object_.find(key);
    return it == object_.end() ? g_null : it->second;
}

const Json &Json::at(size_t i) const
{
    if (type_ != Array || i >= array_.size())
        throw std::runtime_error("json: array index out of range");
    return array_[i];
}

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
    return string_;
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
            v.string_ = parse_string();
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
        i_++;
        skip_ws();
        if (i_ < s_.size() && s_[i_] == ']') {
            i_++;
            return v;
        }
        while (true) {
            v.array_.push_back(parse_value());
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
            v.object_[key] = parse_value();
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
