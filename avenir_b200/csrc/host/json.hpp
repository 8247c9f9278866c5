// Minimal JSON reader (schema files, decision-path lists) -- object member order is preserved.
#pragma once
#include <cstdlib>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace avt_host {

struct Json {
    enum Kind { Null, Bool, Num, Str, Arr, Obj } kind = Null;
    bool b = false;
    double num = 0;
    std::string str;  // Str: value; Num: the literal text
    std::vector<Json> arr;
    std::vector<std::pair<std::string, Json>> obj;

    const Json* get(const std::string& k) const {
        for (auto& kv : obj)
            if (kv.first == k) return &kv.second;
        return nullptr;
    }
    bool is_null() const { return kind == Null; }
};

class JsonParser {
    const std::string& s_;
    size_t i_ = 0;
    [[noreturn]] void err(const char* m) const { throw std::runtime_error(std::string("JSON parse error: ") + m + " at offset " + std::to_string(i_)); }
    void ws() {
        while (i_ < s_.size() && (s_[i_] == ' ' || s_[i_] == '\t' || s_[i_] == '\n' || s_[i_] == '\r')) ++i_;
    }
    static void put_utf8(std::string& o, unsigned cp) {
        if (cp < 0x80) o.push_back((char)cp);
        else if (cp < 0x800) { o.push_back((char)(0xC0 | (cp >> 6))); o.push_back((char)(0x80 | (cp & 0x3F))); }
        else { o.push_back((char)(0xE0 | (cp >> 12))); o.push_back((char)(0x80 | ((cp >> 6) & 0x3F))); o.push_back((char)(0x80 | (cp & 0x3F))); }
    }
    std::string string() {
        if (s_[i_] != '"') err("expected string");
        ++i_;
        std::string o;
        while (i_ < s_.size() && s_[i_] != '"') {
            char c = s_[i_++];
            if (c == '\\') {
                if (i_ >= s_.size()) err("bad escape");
                char e = s_[i_++];
                switch (e) {
                    case 'n': o.push_back('\n'); break;
                    case 't': o.push_back('\t'); break;
                    case 'r': o.push_back('\r'); break;
                    case 'b': o.push_back('\b'); break;
                    case 'f': o.push_back('\f'); break;
                    case 'u': {
                        if (i_ + 4 > s_.size()) err("bad \\u escape");
                        put_utf8(o, (unsigned)strtoul(s_.substr(i_, 4).c_str(), nullptr, 16));
                        i_ += 4;
                        break;
                    }
                    default: o.push_back(e);
                }
            } else {
                o.push_back(c);
            }
        }
        if (i_ >= s_.size()) err("unterminated string");
        ++i_;
        return o;
    }
    Json value() {
        ws();
        if (i_ >= s_.size()) err("unexpected end");
        Json j;
        char c = s_[i_];
        if (c == '{') {
            j.kind = Json::Obj;
            ++i_;
            ws();
            if (s_[i_] == '}') { ++i_; return j; }
            for (;;) {
                ws();
                std::string k = string();
                ws();
                if (s_[i_] != ':') err("expected ':'");
                ++i_;
                j.obj.emplace_back(std::move(k), value());
                ws();
                if (s_[i_] == ',') { ++i_; continue; }
                if (s_[i_] == '}') { ++i_; break; }
                err("expected ',' or '}'");
            }
        } else if (c == '[') {
            j.kind = Json::Arr;
            ++i_;
            ws();
            if (s_[i_] == ']') { ++i_; return j; }
            for (;;) {
                j.arr.push_back(value());
                ws();
                if (s_[i_] == ',') { ++i_; continue; }
                if (s_[i_] == ']') { ++i_; break; }
                err("expected ',' or ']'");
            }
        } else if (c == '"') {
            j.kind = Json::Str;
            j.str = string();
        } else if (s_.compare(i_, 4, "true") == 0) { j.kind = Json::Bool; j.b = true; i_ += 4; }
        else if (s_.compare(i_, 5, "false") == 0) { j.kind = Json::Bool; j.b = false; i_ += 5; }
        else if (s_.compare(i_, 4, "null") == 0) { i_ += 4; }
        else if (s_.compare(i_, 3, "NaN") == 0) { j.kind = Json::Num; j.str = "NaN"; j.num = strtod("nan", nullptr); i_ += 3; }
        else {
            size_t st = i_;
            while (i_ < s_.size() && (isdigit((unsigned char)s_[i_]) || s_[i_] == '-' || s_[i_] == '+' || s_[i_] == '.' || s_[i_] == 'e' || s_[i_] == 'E')) ++i_;
            if (st == i_) err("unexpected character");
            j.kind = Json::Num;
            j.str = s_.substr(st, i_ - st);
            j.num = strtod(j.str.c_str(), nullptr);
        }
        return j;
    }

public:
    explicit JsonParser(const std::string& s) : s_(s) {}
    Json parse() {
        Json j = value();
        ws();
        if (i_ != s_.size()) err("trailing characters");
        return j;
    }
};

inline std::string json_escape(const std::string& s) {
    std::string o = "\"";
    for (char c : s) {
        switch (c) {
            case '"': o += "\\\""; break;
            case '\\': o += "\\\\"; break;
            case '\n': o += "\\n"; break;
            case '\t': o += "\\t"; break;
            case '\r': o += "\\r"; break;
            default:
                if ((unsigned char)c < 0x20) { char b[8]; snprintf(b, sizeof b, "\\u%04x", c); o += b; }
                else o.push_back(c);
        }
    }
    return o + "\"";
}

}  // namespace avt_host
