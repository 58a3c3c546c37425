// Minimal XML reader (elements + attributes; comments, declarations and text are skipped).
// Enough for the MJCF subset the shipped TrajOptKP tasks use.  Host-only.
#pragma once
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace tokp {

struct XmlNode {
  std::string tag;
  std::vector<std::pair<std::string, std::string>> attrs;
  std::vector<std::unique_ptr<XmlNode>> children;

  const std::string* attr(const std::string& k) const {
    for (auto& a : attrs)
      if (a.first == k) return &a.second;
    return nullptr;
  }
  bool has(const std::string& k) const { return attr(k) != nullptr; }
};

class XmlParser {
 public:
  explicit XmlParser(const std::string& text) : s_(text), p_(0) {}

  std::unique_ptr<XmlNode> parse() {
    skip_misc();
    auto root = parse_element();
    if (!root) throw std::runtime_error("xml: no root element");
    return root;
  }

 private:
  const std::string& s_;
  size_t p_;

  bool starts(const char* lit) const { return s_.compare(p_, std::char_traits<char>::length(lit), lit) == 0; }
  void skip_ws() {
    while (p_ < s_.size() && (s_[p_] == ' ' || s_[p_] == '\t' || s_[p_] == '\n' || s_[p_] == '\r')) p_++;
  }
  // skip whitespace, comments, <?...?>, <!DOCTYPE...>, and character data
  void skip_misc() {
    for (;;) {
      while (p_ < s_.size() && s_[p_] != '<') p_++;
      if (p_ >= s_.size()) return;
      if (starts("<!--")) {
        size_t e = s_.find("-->", p_ + 4);
        if (e == std::string::npos) throw std::runtime_error("xml: unterminated comment");
        p_ = e + 3;
      } else if (starts("<?")) {
        size_t e = s_.find("?>", p_ + 2);
        if (e == std::string::npos) throw std::runtime_error("xml: unterminated declaration");
        p_ = e + 2;
      } else if (starts("<!")) {
        size_t e = s_.find('>', p_ + 2);
        if (e == std::string::npos) throw std::runtime_error("xml: unterminated <!");
        p_ = e + 1;
      } else {
        return;
      }
    }
  }
  std::string parse_name() {
    size_t b = p_;
    while (p_ < s_.size()) {
      char c = s_[p_];
      if (c == ' ' || c == '\t' || c == '\n' || c == '\r' || c == '=' || c == '>' || c == '/') break;
      p_++;
    }
    return s_.substr(b, p_ - b);
  }
  std::unique_ptr<XmlNode> parse_element() {
    if (p_ >= s_.size() || s_[p_] != '<') return nullptr;
    p_++;
    auto node = std::make_unique<XmlNode>();
    node->tag = parse_name();
    for (;;) {
      skip_ws();
      if (p_ >= s_.size()) throw std::runtime_error("xml: unexpected end in <" + node->tag);
      if (s_[p_] == '/') {
        if (p_ + 1 >= s_.size() || s_[p_ + 1] != '>') throw std::runtime_error("xml: bad '/>'");
        p_ += 2;
        return node;
      }
      if (s_[p_] == '>') {
        p_++;
        break;
      }
      std::string k = parse_name();
      skip_ws();
      if (p_ >= s_.size() || s_[p_] != '=') throw std::runtime_error("xml: attribute without '=' in <" + node->tag);
      p_++;
      skip_ws();
      char q = s_[p_];
      if (q != '"' && q != '\'') throw std::runtime_error("xml: unquoted attribute in <" + node->tag);
      size_t e = s_.find(q, p_ + 1);
      if (e == std::string::npos) throw std::runtime_error("xml: unterminated attribute value");
      node->attrs.emplace_back(k, s_.substr(p_ + 1, e - p_ - 1));
      p_ = e + 1;
    }
    // children until closing tag
    for (;;) {
      skip_misc();
      if (p_ >= s_.size()) throw std::runtime_error("xml: missing </" + node->tag + ">");
      if (starts("</")) {
        p_ += 2;
        std::string name = parse_name();
        if (name != node->tag) throw std::runtime_error("xml: mismatched </" + name + "> for <" + node->tag + ">");
        skip_ws();
        if (p_ >= s_.size() || s_[p_] != '>') throw std::runtime_error("xml: bad closing tag");
        p_++;
        return node;
      }
      node->children.push_back(parse_element());
    }
  }
};

}  // namespace tokp
