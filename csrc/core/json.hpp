// Minimal JSON value + recursive-descent parser (objects, arrays, strings,
// numbers, bools, null). Enough for the cluster config / stage documents the
// config server and runners exchange.
#pragma once
#include <cstdlib>
#include <map>
#include <string>
#include <vector>

namespace kf {

struct Json {
    enum Kind { Null, Bool, Num, Str, Arr, Obj } kind = Null;
    bool b        = false;
    double num    = 0;
    std::string str;
    std::vector<Json> arr;
    std::map<std::string, Json> obj;

    const Json *get(const std::string &k) const
    {
        auto it = obj.find(k);
        return it == obj.end() ? nullptr : &it->second;
    }
};

class JsonParser
{
    const std::string &s_;
    size_t i_ = 0;
    bool ok_  = true;

    void ws()
    {
        while (i_ < s_.size() && (s_[i_] == ' ' || s_[i_] == '\n' || s_[i_] == '\t' || s_[i_] == '\r')) ++i_;
    }
    bool eat(char c)
    {
        ws();
        if (i_ < s_.size() && s_[i_] == c) {
            ++i_;
            return true;
        }
        return false;
    }
    std::string parse_string()
    {
        std::string out;
        if (!eat('"')) {
            ok_ = false;
            return out;
        }
        while (i_ < s_.size() && s_[i_] != '"') {
            char c = s_[i_++];
            if (c == '\\' && i_ < s_.size()) {
                char e = s_[i_++];
                switch (e) {
                case 'n': out += '\n'; break;
                case 't': out += '\t'; break;
                case 'r': out += '\r'; break;
                case 'b': out += '\b'; break;
                case 'f': out += '\f'; break;
                case 'u':
                    if (i_ + 4 <= s_.size()) {
                        unsigned v = (unsigned)strtoul(s_.substr(i_, 4).c_str(), nullptr, 16);
                        i_ += 4;
                        out += (char)(v < 0x80 ? v : '?');
                    }
                    break;
                default: out += e;
                }
            } else {
                out += c;
            }
        }
        if (i_ >= s_.size()) ok_ = false;
        ++i_;
        return out;
    }
    // Cluster and stage documents nest three levels deep. The limit keeps a hostile "[[[[..." (a 100 KB PUT to the
    // config server is enough) from overflowing the stack of this recursive-descent parser.
    static constexpr int kMaxDepth = 64;
    int depth_ = 0;
    struct DepthGuard {
        int &d;
        explicit DepthGuard(int &x) : d(x) { ++d; }
        ~DepthGuard() { --d; }
    };

    Json value()
    {
        Json j;
        DepthGuard guard(depth_);
        ws();
        if (i_ >= s_.size() || depth_ > kMaxDepth) {
            ok_ = false;
            return j;
        }
        const char c = s_[i_];
        if (c == '{') {
            ++i_;
            j.kind = Json::Obj;
            if (eat('}')) return j;
            do {
                std::string k = parse_string();
                if (!eat(':')) ok_ = false;
                if (!ok_) return j;
                j.obj[k] = value();
            } while (ok_ && eat(','));
            if (!eat('}')) ok_ = false;
        } else if (c == '[') {
            ++i_;
            j.kind = Json::Arr;
            if (eat(']')) return j;
            do {
                j.arr.push_back(value());
            } while (ok_ && eat(','));
            if (!eat(']')) ok_ = false;
        } else if (c == '"') {
            j.kind = Json::Str;
            j.str  = parse_string();
        } else if (s_.compare(i_, 4, "true") == 0) {
            j.kind = Json::Bool;
            j.b    = true;
            i_ += 4;
        } else if (s_.compare(i_, 5, "false") == 0) {
            j.kind = Json::Bool;
            i_ += 5;
        } else if (s_.compare(i_, 4, "null") == 0) {
            i_ += 4;
        } else {
            char *end = nullptr;
            j.num     = strtod(s_.c_str() + i_, &end);
            if (end == s_.c_str() + i_) {
                ok_ = false;
                return j;
            }
            j.kind = Json::Num;
            i_     = end - s_.c_str();
        }
        return j;
    }

  public:
    explicit JsonParser(const std::string &s) : s_(s) {}
    bool parse(Json *out)
    {
        *out = value();
        ws();
        return ok_ && i_ == s_.size();
    }
};

inline std::string json_quote(const std::string &s)
{
    std::string o = "\"";
    for (char c : s) {
        if (c == '"' || c == '\\') {
            o += '\\';
            o += c;
        } else if (c == '\n') {
            o += "\\n";
        } else {
            o += c;
        }
    }
    return o + "\"";
}

}  // namespace kf
