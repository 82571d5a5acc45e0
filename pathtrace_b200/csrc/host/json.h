// json.h -- minimal JSON reader for config.json and the scene files.
// The reference reads both with nlohmann::json 3.7.3 (main.cpp:110-134); only the subset of its
// behaviour the two formats rely on is provided here: objects, arrays, numbers, strings, booleans,
// null; `contains`, `value(key, default)`, `at(i)`, numeric conversion of ints to floats and back.
#pragma once
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace b200pt {

class Json {
public:
    enum Type { Null, Bool, Number, String, Array, Object };

    Json() : type_(Null) {}
    static Json parse(const std::string &text);

    Type type() const { return type_; }
    bool is_array() const { return type_ == Array; }
    bool is_object() const { return type_ == Object; }
    bool is_number() const { return type_ == Number; }
    bool is_string() const { return type_ == String; }
    bool is_null() const { return type_ == Null; }

    bool contains(const std::string &key) const { return type_ == Object && object_.count(key) > 0; }
    // operator[] on a missing key yields null (nlohmann would insert null on a non-const json).
    const Json &operator[](const std::string &key) const;
    const Json &at(size_t i) const;
    size_t size() const { return type_ == Array ? array_.size() : (type_ == Object ? object_.size() : 0); }
    const std::vector<Json> &items() const { return array_; }

    double as_double() const;
    float as_float() const { return (float)as_double(); }
    long as_int() const { return (long)as_double(); }
    bool as_bool() const;
    const std::string &as_string() const;

    double value(const std::string &key, double dflt) const { return contains(key) ? (*this)[key].as_double() : dflt; }
    long value(const std::string &key, long dflt) const { return contains(key) ? (*this)[key].as_int() : dflt; }
    long value(const std::string &key, int dflt) const { return value(key, (long)dflt); }
    bool value(const std::string &key, bool dflt) const { return contains(key) ? (*this)[key].as_bool() : dflt; }
    std::string value(const std::string &key, const char *dflt) const
    {
        return contains(key) ? (*this)[key].as_string() : std::string(dflt);
    }

private:
    Type type_;
    bool bool_ = false;
    double number_ = 0.0;
    std::string string_;
    std::vector<Json> array_;
    std::map<std::string, Json> object_;
    friend class JsonParser;
};

} // namespace b200pt
