// yaml_lite.hpp — reader for the YAML subset the reference's settings files use
// (reference tutorials/*/settings.yml, read there through fkYAML in src/settings.cpp:18-84):
// block mappings nested by indentation, flow sequences of scalars `[a, b, c]`, plain / quoted
// scalars, `#` comments, blank lines. Keys may contain spaces ("particle distance"). Anything else
// (anchors, block sequences, multi-line scalars, multiple documents) is rejected with an error that
// names the line. Own code; no third-party parser.
#pragma once

#include <cstdlib>
#include <fstream>
#include <map>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace yaml_lite {

class Node {
public:
    enum Kind { Null, Scalar, Sequence, Mapping };
    Kind kind = Null;
    std::string scalar;
    std::vector<Node> items;
    std::vector<std::pair<std::string, Node>> entries; // insertion order kept
    std::string path;                                  // for error messages

    bool contains(const std::string& key) const {
        for (auto& e : entries)
            if (e.first == key) return true;
        return false;
    }
    const Node& operator[](const std::string& key) const {
        if (kind != Mapping) throw std::runtime_error("yaml: '" + path + "' is not a mapping (looking up '" + key + "')");
        for (auto& e : entries)
            if (e.first == key) return e.second;
        throw std::runtime_error("yaml: key '" + key + "' not found in '" + (path.empty() ? "<root>" : path) + "'");
    }
    const Node& operator[](std::size_t i) const {
        if (kind != Sequence) throw std::runtime_error("yaml: '" + path + "' is not a sequence");
        if (i >= items.size()) throw std::runtime_error("yaml: index out of range in '" + path + "'");
        return items[i];
    }
    // The reference's files say: "Numbers read as a double in the program must be written to one
    // decimal place" (fkYAML types 1 as an integer and refuses get_value<double>). Here an integer
    // literal is accepted where a double is asked for; the value is the same.
    double asDouble() const {
        need(Scalar);
        char* end    = nullptr;
        double v     = std::strtod(scalar.c_str(), &end);
        if (end == scalar.c_str() || *end != '\0') throw std::runtime_error("yaml: '" + path + "': not a number: " + scalar);
        return v;
    }
    int asInt() const {
        need(Scalar);
        char* end = nullptr;
        long v    = std::strtol(scalar.c_str(), &end, 10);
        if (end == scalar.c_str() || *end != '\0') throw std::runtime_error("yaml: '" + path + "': not an integer: " + scalar);
        return (int) v;
    }
    bool asBool() const {
        need(Scalar);
        if (scalar == "true" || scalar == "True" || scalar == "TRUE") return true;
        if (scalar == "false" || scalar == "False" || scalar == "FALSE") return false;
        throw std::runtime_error("yaml: '" + path + "': not a boolean: " + scalar);
    }
    const std::string& asString() const {
        need(Scalar);
        return scalar;
    }

private:
    void need(Kind k) const {
        if (kind != k) throw std::runtime_error("yaml: '" + path + "' has the wrong kind of value");
    }
};

namespace detail {
inline std::string trim(const std::string& s) {
    std::size_t a = s.find_first_not_of(" \t\r\n"), b = s.find_last_not_of(" \t\r\n");
    return a == std::string::npos ? std::string() : s.substr(a, b - a + 1);
}
// strip a trailing comment: '#' at line start or preceded by whitespace, outside quotes
inline std::string stripComment(const std::string& s) {
    char quote = 0;
    for (std::size_t i = 0; i < s.size(); i++) {
        char c = s[i];
        if (quote) {
            if (c == quote) quote = 0;
        } else if (c == '"' || c == '\'') {
            quote = c;
        } else if (c == '#' && (i == 0 || s[i - 1] == ' ' || s[i - 1] == '\t')) {
            return s.substr(0, i);
        }
    }
    return s;
}
inline std::string unquote(const std::string& s) {
    if (s.size() >= 2 && ((s.front() == '"' && s.back() == '"') || (s.front() == '\'' && s.back() == '\''))) return s.substr(1, s.size() - 2);
    return s;
}
inline Node parseValue(const std::string& text, const std::string& path, int lineNo) {
    Node n;
    n.path = path;
    if (!text.empty() && text.front() == '[') {
        if (text.back() != ']') throw std::runtime_error("yaml: line " + std::to_string(lineNo) + ": unterminated flow sequence");
        n.kind = Node::Sequence;
        std::string inner = text.substr(1, text.size() - 2), item;
        std::stringstream ss(inner);
        int k = 0;
        while (std::getline(ss, item, ',')) {
            Node c;
            c.kind   = Node::Scalar;
            c.scalar = unquote(trim(item));
            c.path   = path + "[" + std::to_string(k++) + "]";
            n.items.push_back(c);
        }
        return n;
    }
    if (!text.empty() && (text.front() == '&' || text.front() == '*' || text.front() == '|' || text.front() == '>' || text.front() == '{'))
        throw std::runtime_error("yaml: line " + std::to_string(lineNo) + ": unsupported YAML construct: " + text);
    n.kind   = Node::Scalar;
    n.scalar = unquote(text);
    return n;
}
} // namespace detail

inline Node parse(std::istream& in) {
    struct Frame {
        int indent;
        Node* node;
    };
    Node root;
    root.kind = Node::Mapping;
    std::vector<Frame> stack{{-1, &root}};
    // pointers into `entries` vectors stay valid only while no sibling is appended; a frame's node is
    // the LAST entry of its parent, and a parent only grows after its child frames are popped.
    std::string raw;
    int lineNo = 0;
    while (std::getline(in, raw)) {
        lineNo++;
        std::string line = detail::stripComment(raw);
        if (detail::trim(line).empty()) continue;
        if (line.find('\t') != std::string::npos && line.find_first_not_of(" \t") > line.find('\t'))
            throw std::runtime_error("yaml: line " + std::to_string(lineNo) + ": tab in indentation");
        int indent       = (int) line.find_first_not_of(' ');
        std::string body = detail::trim(line);
        if (body == "---" || body == "...") continue;
        if (body.front() == '-') throw std::runtime_error("yaml: line " + std::to_string(lineNo) + ": block sequences are not supported");
        std::size_t colon = std::string::npos;
        {
            char quote = 0;
            for (std::size_t i = 0; i < body.size(); i++) {
                char c = body[i];
                if (quote) {
                    if (c == quote) quote = 0;
                } else if (c == '"' || c == '\'') {
                    quote = c;
                } else if (c == ':' && (i + 1 == body.size() || body[i + 1] == ' ')) {
                    colon = i;
                    break;
                }
            }
        }
        if (colon == std::string::npos) throw std::runtime_error("yaml: line " + std::to_string(lineNo) + ": expected 'key: value'");
        std::string key  = detail::unquote(detail::trim(body.substr(0, colon)));
        std::string rest = detail::trim(body.substr(colon + 1));
        while (stack.size() > 1 && indent <= stack.back().indent) stack.pop_back();
        Node* parent = stack.back().node;
        if (parent->kind != Node::Mapping) throw std::runtime_error("yaml: line " + std::to_string(lineNo) + ": bad indentation");
        if (parent->contains(key)) throw std::runtime_error("yaml: line " + std::to_string(lineNo) + ": duplicate key '" + key + "'");
        std::string path = parent->path.empty() ? key : parent->path + "." + key;
        if (rest.empty()) {
            Node m;
            m.kind = Node::Mapping;
            m.path = path;
            parent->entries.emplace_back(key, m);
            stack.push_back({indent, &parent->entries.back().second});
        } else {
            parent->entries.emplace_back(key, detail::parseValue(rest, path, lineNo));
        }
    }
    return root;
}

inline Node parseFile(const std::string& file) {
    std::ifstream ifs(file);
    if (!ifs.is_open()) throw std::runtime_error("yaml: cannot open " + file);
    return parse(ifs);
}

} // namespace yaml_lite
