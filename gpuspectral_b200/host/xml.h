// xml.h — a small XML reader, enough for Mitsuba scene files: elements, attributes,
// comments, declarations.  Text content is skipped (Mitsuba scenes keep all data in attributes).
#pragma once

#include <memory>
#include <string>
#include <utility>
#include <vector>

namespace gsbhost {

struct XmlNode {
    std::string tag;
    std::vector<std::pair<std::string, std::string>> attrs;
    std::vector<std::unique_ptr<XmlNode>> children;

    const std::string* attr(const std::string& name) const {
        for (auto& a : attrs)
            if (a.first == name) return &a.second;
        return nullptr;
    }
    std::string attrOr(const std::string& name, const std::string& fallback) const {
        const std::string* v = attr(name);
        return v ? *v : fallback;
    }
};

// Parses a whole document and returns its root element. Throws std::runtime_error on malformed input.
std::unique_ptr<XmlNode> parseXml(const std::string& text);

}  // namespace gsbhost
