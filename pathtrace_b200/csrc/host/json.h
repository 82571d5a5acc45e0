// json.h -- minimal JSON reader for config.json and the scene files.
// The reference reads both with nlohmann::json 3.7.3 (main.cpp:110-134); only the subset of its
// behaviour the two formats rely on is provided here: objects, arrays, numbers, strings, booleans,
// null; `contains`, `value(key, default)`, `at(i)`, numeric conversion of ints to floats and back.
#pragma once
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace b200pt {

// A value is 40 bytes: a number lives in the value itself, strings / arrays / objects behind a pointer.  Objects keep
// their members in file order and are searched linearly (the two formats have at most a dozen keys per object); a
// scene file of 8192 instances is ~100 000 values.
class Json {
public:
    enum Type { Null, Bool, Number, String, Array, Object };
    using Members = std::vector<std::pair<std::string, Json>>;

    Json() = default;
    Json(Json &&) = default;
    Json &operator=(Json &&) = default;
    Json(const Json &) = delete;
    Json &operator=(const Json &) = delete;
    static Json parse(const std::string &text);

    Type type() const { return type_; }
    bool is_array() const { return type_ == Array; }
    bool is_object() const { return type_ == Object; }
    bool is_number() const { return type_ == Number; }
    bool is_string() const { return type_ == String; }
    bool is_null() const { return type_ == Null; }

    bool contains(const char *key) const { return find(key) != nullptr; }
    bool contains(const std::string &key) const { return find(key.c_str()) != nullptr; }
    // operator[] on a missing key yields null (nlohmann would insert null on a non-const json).
    const Json &operator[](const char *key) const;
    const Json &operator[](const std::string &key) const { return (*this)[key.c_str()]; }
    const Json &at(size_t i) const;
    size_t size() const { return type_ == Array ? array_->size() : (type_ == Object ? object_->size() : 0); }
    const std::vector<Json> &items() const;

    double as_double() const;
    float as_float() const { return (float)as_double(); }
    long as_int() const { return (long)as_double(); }
    bool as_bool() const;
    const std::string &as_string() const;

    double value(const char *key, double dflt) const
    {
        const Json *j = find(key);
        return j ? j->as_double() : dflt;
    }
    long value(const char *key, long dflt) const
    {
        const Json *j = find(key);
        return j ? j->as_int() : dflt;
    }
    long value(const char *key, int dflt) const { return value(key, (long)dflt); }
    bool value(const char *key, bool dflt) const
    {
        const Json *j = find(key);
        return j ? j->as_bool() : dflt;
    }
    std::string value(const char *key, const char *dflt) const
    {
        const Json *j = find(key);
        return j ? j->as_string() : std::string(dflt);
    }

private:
    const Json *find(const char *key) const;

    Type type_ = Null;
    bool bool_ = false;
    double number_ = 0.0;
    std::unique_ptr<std::string> string_;
    std::unique_ptr<std::vector<Json>> array_;
    std::unique_ptr<Members> object_;
    friend class JsonParser;
};

} // namespace b200pt
