// rt_yaml.h — small YAML reader for scene files (header-only, no dependencies).
//
// The reference loads scenes with yaml-cpp (src/scene.cpp:3,154-203), which is not on this image.  This reader
// covers the subset scene files use — block mappings and sequences, flow sequences / mappings ([a, b], {k: v}),
// plain and quoted scalars, comments, "---" — and keeps yaml-cpp's Mark (0-based line / column of every node) so
// that the loader's error messages carry the same "line: L column: C" positions (scene.cpp:24-29).
#ifndef RT_YAML_H
#define RT_YAML_H

#include <cctype>
#include <cerrno>
#include <cmath>
#include <cstdlib>
#include <fstream>
#include <istream>
#include <sstream>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace rt_yaml {

struct Mark {
    int line = 0, column = 0;   // 0-based, like YAML::Mark
};

struct ParseError : std::runtime_error {
    Mark mark;
    ParseError(const std::string &what, Mark m)
        : std::runtime_error("yaml-cpp: error at line " + std::to_string(m.line + 1) + ", column " +
                             std::to_string(m.column + 1) + ": " + what), mark(m) {}
};
struct BadFile : std::runtime_error {
    explicit BadFile(const std::string &path) : std::runtime_error("bad file: " + path) {}
};
struct BadConversion : std::runtime_error {
    Mark mark;
    explicit BadConversion(Mark m) : std::runtime_error("bad conversion"), mark(m) {}
};

class Node {
public:
    enum Type { Undefined, Null, Scalar, Sequence, Map };
    Type type = Undefined;
    Mark mark;
    std::string scalar;
    std::vector<Node> items;                               // Sequence
    std::vector<std::pair<std::string, Node>> entries;     // Map, in file order

    bool IsDefined() const { return type != Undefined; }
    bool IsSequence() const { return type == Sequence; }
    bool IsMap() const { return type == Map; }
    bool IsScalar() const { return type == Scalar; }
    size_t size() const { return type == Sequence ? items.size() : (type == Map ? entries.size() : 0); }

    // map lookup; an undefined node (carrying the parent's mark) when absent — like yaml-cpp's const operator[]
    const Node &operator[](const char *key) const
    {
        if (type == Map)
            for (const auto &e : entries)
                if (e.first == key) return e.second;
        return undefined();
    }
    const Node &operator[](size_t i) const { return (type == Sequence && i < items.size()) ? items[i] : undefined(); }

    // conversions: throw BadConversion(mark) like node.as<T>()
    double as_double() const
    {
        if (type != Scalar) throw BadConversion(mark);
        const std::string &s = scalar;
        if (s == ".inf" || s == ".Inf" || s == ".INF" || s == "+.inf" || s == "+.Inf" || s == "+.INF") return INFINITY;
        if (s == "-.inf" || s == "-.Inf" || s == "-.INF") return -INFINITY;
        if (s == ".nan" || s == ".NaN" || s == ".NAN") return NAN;
        if (s.empty()) throw BadConversion(mark);
        errno = 0;
        char *end = nullptr;
        const double v = std::strtod(s.c_str(), &end);
        if (end == s.c_str() || *end != '\0') throw BadConversion(mark);
        // strtod accepts "inf"/"nan"/hex floats, an iostream extraction (what yaml-cpp uses) does not
        for (char ch : s)
            if (!(std::isdigit((unsigned char) ch) || ch == '+' || ch == '-' || ch == '.' || ch == 'e' || ch == 'E'))
                throw BadConversion(mark);
        return v;
    }
    float as_float() const { return (float) as_double(); }
    unsigned as_unsigned() const
    {
        if (type != Scalar || scalar.empty() || scalar[0] == '-') throw BadConversion(mark);
        errno = 0;
        char *end = nullptr;
        const unsigned long long v = std::strtoull(scalar.c_str(), &end, 0);
        if (end == scalar.c_str() || *end != '\0' || errno == ERANGE || v > 0xffffffffULL) throw BadConversion(mark);
        return (unsigned) v;
    }
    const std::string &as_string() const
    {
        if (type != Scalar) throw BadConversion(mark);
        return scalar;
    }

private:
    const Node &undefined() const
    {
        static thread_local Node u;
        u = Node();
        u.mark = mark;
        return u;
    }
};

namespace detail {

struct Line {
    int indent;
    int number;            // 0-based line in the file
    std::string text;      // content without indentation, comment and trailing blanks
};

inline std::string rstrip(std::string s)
{
    while (!s.empty() && (s.back() == ' ' || s.back() == '\t' || s.back() == '\r')) s.pop_back();
    return s;
}

// cut a trailing comment: '#' at the start or after a blank, outside quotes
inline std::string strip_comment(const std::string &s)
{
    char quote = 0;
    for (size_t i = 0; i < s.size(); i++) {
        const char ch = s[i];
        if (quote) {
            if (ch == quote) quote = 0;
        }
        else if (ch == '"' || ch == '\'') quote = ch;
        else if (ch == '#' && (i == 0 || s[i - 1] == ' ' || s[i - 1] == '\t')) return s.substr(0, i);
    }
    return s;
}

inline std::string unquote(const std::string &s, Mark m)
{
    if (s.size() >= 2 && (s.front() == '"' || s.front() == '\'')) {
        if (s.back() != s.front()) throw ParseError("end of quoted scalar not found", m);
        return s.substr(1, s.size() - 2);
    }
    return s;
}

// position of the ':' that ends a mapping key in `s` (followed by a blank or the end), outside quotes / brackets
inline size_t key_colon(const std::string &s)
{
    char quote = 0;
    int depth = 0;
    for (size_t i = 0; i < s.size(); i++) {
        const char ch = s[i];
        if (quote) {
            if (ch == quote) quote = 0;
            continue;
        }
        if (ch == '"' || ch == '\'') quote = ch;
        else if (ch == '[' || ch == '{') depth++;
        else if (ch == ']' || ch == '}') depth--;
        else if (ch == ':' && depth == 0 && (i + 1 == s.size() || s[i + 1] == ' ' || s[i + 1] == '\t')) return i;
    }
    return std::string::npos;
}

class Parser {
public:
    explicit Parser(std::vector<Line> lines) : lines_(std::move(lines)) {}

    Node document()
    {
        if (lines_.empty()) {
            Node n;
            n.type = Node::Null;
            return n;
        }
        Node n = block(lines_[0].indent);
        if (pos_ < lines_.size()) throw ParseError("unexpected content", Mark{lines_[pos_].number, lines_[pos_].indent});
        return n;
    }

private:
    std::vector<Line> lines_;
    size_t pos_ = 0;

    static bool is_item(const std::string &t) { return t == "-" || (t.size() > 1 && t[0] == '-' && (t[1] == ' ' || t[1] == '\t')); }

    // a node written inline after "key:" or "- ": flow collection or scalar
    Node inline_value(const std::string &text, Mark m)
    {
        size_t p = 0;
        Node n = flow(text, p, m);
        while (p < text.size() && (text[p] == ' ' || text[p] == '\t')) p++;
        if (p != text.size()) throw ParseError("unexpected characters after value", Mark{m.line, m.column + (int) p});
        return n;
    }

    Node flow(const std::string &s, size_t &p, Mark base)
    {
        while (p < s.size() && (s[p] == ' ' || s[p] == '\t')) p++;
        Node n;
        n.mark = Mark{base.line, base.column + (int) p};
        if (p < s.size() && (s[p] == '[' || s[p] == '{')) {
            const bool seq = s[p] == '[';
            const char close = seq ? ']' : '}';
            n.type = seq ? Node::Sequence : Node::Map;
            p++;
            for (;;) {
                while (p < s.size() && (s[p] == ' ' || s[p] == '\t')) p++;
                if (p >= s.size()) throw ParseError(seq ? "end of sequence flow not found" : "end of map flow not found", n.mark);
                if (s[p] == close) {
                    p++;
                    break;
                }
                if (seq) {
                    n.items.push_back(flow(s, p, base));
                }
                else {
                    size_t k0 = p;
                    while (p < s.size() && s[p] != ':' && s[p] != ',' && s[p] != '}') p++;
                    if (p >= s.size() || s[p] != ':') throw ParseError("illegal map value", Mark{base.line, base.column + (int) k0});
                    std::string key = unquote(rstrip(s.substr(k0, p - k0)), n.mark);
                    p++;
                    n.entries.emplace_back(key, flow(s, p, base));
                }
                while (p < s.size() && (s[p] == ' ' || s[p] == '\t')) p++;
                if (p < s.size() && s[p] == ',') p++;
                else if (p < s.size() && s[p] != close) throw ParseError("illegal flow end", Mark{base.line, base.column + (int) p});
            }
            return n;
        }
        // scalar: quoted, or plain up to a flow delimiter
        size_t q = p;
        if (q < s.size() && (s[q] == '"' || s[q] == '\'')) {
            const char quote = s[q++];
            while (q < s.size() && s[q] != quote) q++;
            if (q >= s.size()) throw ParseError("end of quoted scalar not found", n.mark);
            n.scalar = s.substr(p + 1, q - p - 1);
            p = q + 1;
            n.type = Node::Scalar;
            return n;
        }
        while (q < s.size() && s[q] != ',' && s[q] != ']' && s[q] != '}') q++;
        n.scalar = rstrip(s.substr(p, q - p));
        p = q;
        if (n.scalar.empty() || n.scalar == "~" || n.scalar == "null") n.type = Node::Null;
        else n.type = Node::Scalar;
        return n;
    }

    // a plain scalar may itself contain , ] } when it is not inside a flow collection
    Node plain_or_flow(const std::string &text, Mark m)
    {
        if (!text.empty() && (text[0] == '[' || text[0] == '{' || text[0] == '"' || text[0] == '\'')) return inline_value(text, m);
        Node n;
        n.mark = m;
        n.scalar = text;
        n.type = (text.empty() || text == "~" || text == "null") ? Node::Null : Node::Scalar;
        return n;
    }

    // the value of "key:" or "-" whose text continues on the following, deeper lines (or is absent)
    Node nested(int parent_indent, bool allow_same_indent_sequence, Mark parent)
    {
        if (pos_ < lines_.size()) {
            const Line &nx = lines_[pos_];
            if (nx.indent > parent_indent) return block(nx.indent);
            if (allow_same_indent_sequence && nx.indent == parent_indent && is_item(nx.text)) return block(nx.indent);
        }
        Node n;
        n.type = Node::Null;
        n.mark = parent;
        return n;
    }

    Node block(int indent)
    {
        const Line first = lines_[pos_];
        Node n;
        n.mark = Mark{first.number, indent};
        if (is_item(first.text)) {
            n.type = Node::Sequence;
            while (pos_ < lines_.size() && lines_[pos_].indent == indent && is_item(lines_[pos_].text)) {
                Line &ln = lines_[pos_];
                size_t off = 1;
                while (off < ln.text.size() && (ln.text[off] == ' ' || ln.text[off] == '\t')) off++;
                const std::string rest = ln.text.substr(off);
                const Mark im{ln.number, indent + (int) off};
                if (rest.empty()) {
                    pos_++;
                    n.items.push_back(nested(indent, false, im));
                }
                else if (key_colon(rest) != std::string::npos && rest[0] != '[' && rest[0] != '{') {
                    // "- key: value": the item is a mapping that starts on this line
                    ln.indent = indent + (int) off;
                    ln.text = rest;
                    n.items.push_back(block(ln.indent));
                }
                else {
                    pos_++;
                    n.items.push_back(plain_or_flow(rest, im));
                }
            }
            if (pos_ < lines_.size() && lines_[pos_].indent > indent)
                throw ParseError("bad indentation of a sequence entry", Mark{lines_[pos_].number, lines_[pos_].indent});
            return n;
        }
        if (key_colon(first.text) != std::string::npos && first.text[0] != '[' && first.text[0] != '{') {
            n.type = Node::Map;
            while (pos_ < lines_.size() && lines_[pos_].indent == indent && !is_item(lines_[pos_].text)) {
                const Line ln = lines_[pos_];
                const size_t c = key_colon(ln.text);
                if (c == std::string::npos) throw ParseError("illegal map value", Mark{ln.number, indent});
                const std::string key = unquote(rstrip(ln.text.substr(0, c)), Mark{ln.number, indent});
                size_t v = c + 1;
                while (v < ln.text.size() && (ln.text[v] == ' ' || ln.text[v] == '\t')) v++;
                const std::string rest = ln.text.substr(v);
                pos_++;
                if (rest.empty()) n.entries.emplace_back(key, nested(indent, true, Mark{ln.number, indent + (int) v}));
                else n.entries.emplace_back(key, plain_or_flow(rest, Mark{ln.number, indent + (int) v}));
            }
            if (pos_ < lines_.size() && lines_[pos_].indent > indent)
                throw ParseError("bad indentation of a mapping entry", Mark{lines_[pos_].number, lines_[pos_].indent});
            return n;
        }
        pos_++;
        return plain_or_flow(first.text, n.mark);
    }
};

} // namespace detail

inline Node Load(std::istream &in)
{
    std::vector<detail::Line> lines;
    std::string raw;
    for (int number = 0; std::getline(in, raw); number++) {
        std::string s = detail::rstrip(detail::strip_comment(raw));
        size_t ind = 0;
        while (ind < s.size() && s[ind] == ' ') ind++;
        if (ind < s.size() && s[ind] == '\t') throw ParseError("tabs are not allowed as indentation", Mark{number, (int) ind});
        if (ind == s.size()) continue;
        if (s == "---" && lines.empty()) continue;
        if (s == "..." || s == "---") break;
        lines.push_back(detail::Line{(int) ind, number, s.substr(ind)});
    }
    return detail::Parser(std::move(lines)).document();
}

inline Node LoadFile(const std::string &path)
{
    std::ifstream in(path);
    if (!in) throw BadFile(path);
    return Load(in);
}

} // namespace rt_yaml
#endif
