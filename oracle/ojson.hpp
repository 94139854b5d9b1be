// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product path.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// leg may build, load or execute anything under oracle/.
//
// Minimal JSON reader with the semantics the reference gets from
// Boost.PropertyTree's read_json (object_factory.hpp:741-758,
// polygoniser_settings.hpp:147-163): every scalar is kept as its source text and
// converted on demand; arrays are objects whose keys are empty; dotted paths
// ("box.xmin") walk nested objects; a failed conversion yields the default.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <utility>
#include <vector>

namespace ojson {

struct Node {
    std::string text;                                   // scalar text (numbers, strings, true/false/null)
    std::vector<std::pair<std::string, Node>> kids;     // object members / array elements (empty key)

    const Node* child(const std::string& key) const {
        for (auto& kv : kids)
            if (kv.first == key) return &kv.second;
        return nullptr;
    }
    const Node* path(const std::string& dotted) const {
        const Node* cur = this;
        size_t pos = 0;
        while (cur && pos <= dotted.size()) {
            size_t dot = dotted.find('.', pos);
            std::string key = dotted.substr(pos, dot == std::string::npos ? std::string::npos : dot - pos);
            cur = cur->child(key);
            if (dot == std::string::npos) break;
            pos = dot + 1;
        }
        return cur;
    }
    // stream-extraction-like conversions: whole text must be consumed (boost's
    // stream_translator requires eof after skipping whitespace).
    bool as_float(float* out) const {
        if (text.empty()) return false;
        char* end = nullptr;
        float v = strtof(text.c_str(), &end);
        if (end == text.c_str()) return false;
        while (*end == ' ' || *end == '\t' || *end == '\n' || *end == '\r') ++end;
        if (*end) return false;
        *out = v;
        return true;
    }
    bool as_int(int* out) const {
        if (text.empty()) return false;
        char* end = nullptr;
        long v = strtol(text.c_str(), &end, 10);
        if (end == text.c_str()) return false;
        while (*end == ' ' || *end == '\t' || *end == '\n' || *end == '\r') ++end;
        if (*end) return false;
        *out = (int)v;
        return true;
    }
    bool as_bool(bool* out) const {   // boost: try numeric 0/1 first, then boolalpha
        if (text == "0") { *out = false; return true; }
        if (text == "1") { *out = true; return true; }
        if (text == "true") { *out = true; return true; }
        if (text == "false") { *out = false; return true; }
        return false;
    }
    float get_float(const std::string& p, float dflt) const {
        const Node* n = path(p); float v;
        return (n && n->as_float(&v)) ? v : dflt;
    }
    int get_int(const std::string& p, int dflt) const {
        const Node* n = path(p); int v;
        return (n && n->as_int(&v)) ? v : dflt;
    }
    bool get_bool(const std::string& p, bool dflt) const {
        const Node* n = path(p); bool v;
        return (n && n->as_bool(&v)) ? v : dflt;
    }
};

class Parser {
    const char* s;
    [[noreturn]] void fail(const char* why) {
        fprintf(stderr, "oracle json: parse error: %s near '%.20s'\n", why, s);
        abort();   // the reference dies on an uncaught ptree exception here
    }
    void ws() { while (*s == ' ' || *s == '\t' || *s == '\n' || *s == '\r') ++s; }
    std::string str() {
        std::string out;
        ++s;  // opening quote
        while (*s && *s != '"') {
            if (*s == '\\') {
                ++s;
                switch (*s) {
                    case 'n': out += '\n'; break;
                    case 't': out += '\t'; break;
                    case 'r': out += '\r'; break;
                    case 'b': out += '\b'; break;
                    case 'f': out += '\f'; break;
                    case 'u': {   // keep BMP code points as UTF-8
                        unsigned cp = 0;
                        for (int i = 0; i < 4 && s[1]; i++) { ++s; cp = cp * 16 + (unsigned)strtol(std::string(1, *s).c_str(), nullptr, 16); }
                        if (cp < 0x80) out += (char)cp;
                        else if (cp < 0x800) { out += (char)(0xC0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3F)); }
                        else { out += (char)(0xE0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
                        break;
                    }
                    default: out += *s;
                }
                ++s;
            } else {
                out += *s++;
            }
        }
        if (*s != '"') fail("unterminated string");
        ++s;
        return out;
    }
    Node value() {
        ws();
        Node n;
        if (*s == '{') {
            ++s; ws();
            if (*s == '}') { ++s; return n; }
            for (;;) {
                ws();
                if (*s != '"') fail("expected key");
                std::string k = str();
                ws();
                if (*s != ':') fail("expected ':'");
                ++s;
                Node v = value();
                n.kids.emplace_back(std::move(k), std::move(v));
                ws();
                if (*s == ',') { ++s; continue; }
                if (*s == '}') { ++s; break; }
                fail("expected ',' or '}'");
            }
        } else if (*s == '[') {
            ++s; ws();
            if (*s == ']') { ++s; return n; }
            for (;;) {
                Node v = value();
                n.kids.emplace_back(std::string(), std::move(v));
                ws();
                if (*s == ',') { ++s; continue; }
                if (*s == ']') { ++s; break; }
                fail("expected ',' or ']'");
            }
        } else if (*s == '"') {
            n.text = str();
        } else {
            const char* b = s;
            while (*s && *s != ',' && *s != '}' && *s != ']' && *s != ' ' && *s != '\t' && *s != '\n' && *s != '\r') ++s;
            if (s == b) fail("expected value");
            n.text.assign(b, s);
        }
        return n;
    }
public:
    explicit Parser(const char* src) : s(src) {}
    Node parse() {
        Node n = value();
        ws();
        if (*s) fail("trailing characters");
        return n;
    }
};

inline Node parse(const char* src) { return Parser(src).parse(); }

}  // namespace ojson
