// jinja_lite.cpp — see jinja_lite.h.  Host-only; compiled into libaios_b200.so.
#include "jinja_lite.h"

#include <algorithm>
#include <cctype>
#include <cstdint>
#include <cstring>
#include <map>
#include <memory>
#include <stdexcept>

namespace aios {
namespace {

struct TplError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

// ---------------------------------------------------------------- values
struct Value;
using List = std::vector<Value>;
using Dict = std::map<std::string, Value>;
struct Value {
    enum Kind { UNDEF, NONE, BOOL, INT, STR, LIST, DICT } kind = UNDEF;
    bool b = false;
    int64_t i = 0;
    std::string s;
    std::shared_ptr<List> list;
    std::shared_ptr<Dict> dict;  // dicts are shared: `set ns.x = ...` must be seen through every reference (namespace)
    static Value none() { Value v; v.kind = NONE; return v; }
    static Value boolean(bool x) { Value v; v.kind = BOOL; v.b = x; return v; }
    static Value integer(int64_t x) { Value v; v.kind = INT; v.i = x; return v; }
    static Value str(std::string x) { Value v; v.kind = STR; v.s = std::move(x); return v; }
    static Value mklist(List x) { Value v; v.kind = LIST; v.list = std::make_shared<List>(std::move(x)); return v; }
    static Value mkdict(Dict x) { Value v; v.kind = DICT; v.dict = std::make_shared<Dict>(std::move(x)); return v; }
    bool truthy() const {
        switch (kind) {
            case BOOL: return b;
            case INT: return i != 0;
            case STR: return !s.empty();
            case LIST: return !list->empty();
            case DICT: return !dict->empty();
            default: return false;
        }
    }
    std::string to_str() const {
        switch (kind) {
            case NONE: return "None";
            case BOOL: return b ? "True" : "False";
            case INT: return std::to_string(i);
            case STR: return s;
            case UNDEF: return "";
            default: throw TplError("cannot print a list or a mapping");
        }
    }
};
bool equal(const Value &a, const Value &b) {
    if (a.kind == Value::INT && b.kind == Value::BOOL) return a.i == (b.b ? 1 : 0);
    if (a.kind == Value::BOOL && b.kind == Value::INT) return (a.b ? 1 : 0) == b.i;
    if (a.kind != b.kind) return false;
    switch (a.kind) {
        case Value::BOOL: return a.b == b.b;
        case Value::INT: return a.i == b.i;
        case Value::STR: return a.s == b.s;
        case Value::LIST: {
            if (a.list->size() != b.list->size()) return false;
            for (size_t k = 0; k < a.list->size(); k++)
                if (!equal((*a.list)[k], (*b.list)[k])) return false;
            return true;
        }
        case Value::DICT: return a.dict == b.dict;
        default: return true;  // none == none, undefined == undefined
    }
}

bool is_space(char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r' || c == '\f' || c == '\v'; }
std::string strip(const std::string &s, bool left, bool right) {
    size_t a = 0, b = s.size();
    if (left) while (a < b && is_space(s[a])) a++;
    if (right) while (b > a && is_space(s[b - 1])) b--;
    return s.substr(a, b - a);
}

// ---------------------------------------------------------------- expressions
struct Tok {
    enum Kind { END, NAME, INT, STR, OP } kind = END;
    std::string text;
    int64_t ival = 0;
};
std::vector<Tok> lex_expr(const std::string &src) {
    std::vector<Tok> out;
    size_t i = 0;
    const size_t n = src.size();
    while (i < n) {
        const char c = src[i];
        if (is_space(c)) { i++; continue; }
        Tok t;
        if (std::isalpha((unsigned char)c) || c == '_') {
            size_t j = i;
            while (j < n && (std::isalnum((unsigned char)src[j]) || src[j] == '_')) j++;
            t.kind = Tok::NAME; t.text = src.substr(i, j - i);
            i = j;
        } else if (std::isdigit((unsigned char)c)) {
            size_t j = i;
            while (j < n && std::isdigit((unsigned char)src[j])) j++;
            if (j < n && src[j] == '.' && j + 1 < n && std::isdigit((unsigned char)src[j + 1])) throw TplError("float literals are not supported");
            t.kind = Tok::INT; t.ival = std::stoll(src.substr(i, j - i));
            i = j;
        } else if (c == '\'' || c == '"') {
            size_t j = i + 1;
            std::string s;
            while (j < n && src[j] != c) {
                if (src[j] == '\\' && j + 1 < n) {
                    const char e = src[j + 1];
                    if (e == 'n') s += '\n';
                    else if (e == 't') s += '\t';
                    else if (e == 'r') s += '\r';
                    else if (e == '\\' || e == '\'' || e == '"') s += e;
                    else { s += '\\'; s += e; }
                    j += 2;
                } else {
                    s += src[j++];
                }
            }
            if (j >= n) throw TplError("unterminated string literal");
            t.kind = Tok::STR; t.text = std::move(s);
            i = j + 1;
        } else {
            static const char *two[] = {"==", "!=", "<=", ">=", "//", "**"};
            t.kind = Tok::OP;
            bool found = false;
            for (const char *op : two)
                if (src.compare(i, 2, op) == 0) { t.text = op; i += 2; found = true; break; }
            if (!found) {
                if (!strchr("+-*/%~<>()[]{}.,:|=", c)) throw TplError(std::string("unexpected character '") + c + "'");
                t.text = std::string(1, c);
                i++;
            }
        }
        out.push_back(std::move(t));
    }
    out.push_back(Tok{});
    return out;
}

struct Expr;
using ExprP = std::shared_ptr<Expr>;
struct Expr {
    enum Kind { LIT, NAME, ATTR, INDEX, SLICE, CALL, FILTER, TEST, UNARY, BINARY, COND, LISTLIT, DICTLIT } kind = LIT;
    Value lit;
    std::string name;  // NAME / ATTR / FILTER / TEST / operator / CALL of a bare function
    std::vector<ExprP> kids;
    std::vector<std::pair<std::string, ExprP>> kwargs;  // CALL (namespace(...)) and DICTLIT keys
    bool negate = false;  // TEST: `is not`
};
ExprP mk(Expr::Kind k) { auto e = std::make_shared<Expr>(); e->kind = k; return e; }

class Parser {
   public:
    explicit Parser(const std::string &src) : toks_(lex_expr(src)) {}
    ExprP parse_all() {
        ExprP e = expr();
        if (peek().kind != Tok::END) throw TplError("unexpected '" + peek().text + "' in expression");
        return e;
    }
    // for the statement parser
    const Tok &peek() const { return toks_[pos_]; }
    Tok next() { return toks_[pos_++]; }
    bool is_name(const char *w) const { return peek().kind == Tok::NAME && peek().text == w; }
    bool is_op(const char *w) const { return peek().kind == Tok::OP && peek().text == w; }
    bool eat_name(const char *w) { if (is_name(w)) { pos_++; return true; } return false; }
    bool eat_op(const char *w) { if (is_op(w)) { pos_++; return true; } return false; }
    void need_op(const char *w) { if (!eat_op(w)) throw TplError(std::string("expected '") + w + "'"); }
    ExprP expr() {  // a if cond else b
        ExprP e = or_();
        if (is_name("if")) {
            pos_++;
            ExprP c = or_();
            ExprP alt = mk(Expr::LIT);  // undefined when there is no else
            if (eat_name("else")) alt = expr();
            ExprP r = mk(Expr::COND);
            r->kids = {c, e, alt};
            return r;
        }
        return e;
    }
    ExprP postfix_only() { return postfix(); }

   private:
    ExprP bin(const std::string &op, ExprP a, ExprP b) { ExprP e = mk(Expr::BINARY); e->name = op; e->kids = {a, b}; return e; }
    ExprP or_() {
        ExprP e = and_();
        while (eat_name("or")) e = bin("or", e, and_());
        return e;
    }
    ExprP and_() {
        ExprP e = not_();
        while (eat_name("and")) e = bin("and", e, not_());
        return e;
    }
    ExprP not_() {
        if (eat_name("not")) { ExprP e = mk(Expr::UNARY); e->name = "not"; e->kids = {not_()}; return e; }
        return cmp();
    }
    ExprP cmp() {
        ExprP e = concat();
        for (;;) {
            if (peek().kind == Tok::OP && (peek().text == "==" || peek().text == "!=" || peek().text == "<" || peek().text == ">" ||
                                           peek().text == "<=" || peek().text == ">=")) {
                const std::string op = next().text;
                e = bin(op, e, concat());
            } else if (is_name("in")) {
                pos_++;
                e = bin("in", e, concat());
            } else if (is_name("not") && toks_[pos_ + 1].kind == Tok::NAME && toks_[pos_ + 1].text == "in") {
                pos_ += 2;
                ExprP n = mk(Expr::UNARY); n->name = "not"; n->kids = {bin("in", e, concat())};
                e = n;
            } else if (is_name("is")) {
                pos_++;
                ExprP t = mk(Expr::TEST);
                t->negate = eat_name("not");
                if (peek().kind != Tok::NAME) throw TplError("expected a test name after 'is'");
                t->name = next().text;
                t->kids = {e};
                e = t;
            } else {
                return e;
            }
        }
    }
    ExprP concat() {
        ExprP e = add();
        while (eat_op("~")) e = bin("~", e, add());
        return e;
    }
    ExprP add() {
        ExprP e = mul();
        while (is_op("+") || is_op("-")) { const std::string op = next().text; e = bin(op, e, mul()); }
        return e;
    }
    ExprP mul() {
        ExprP e = unary();
        while (is_op("*") || is_op("/") || is_op("//") || is_op("%")) { const std::string op = next().text; e = bin(op, e, unary()); }
        return e;
    }
    ExprP unary() {
        if (is_op("-") || is_op("+")) { ExprP e = mk(Expr::UNARY); e->name = next().text; e->kids = {unary()}; return e; }
        return filtered();
    }
    ExprP filtered() {  // filters bind tighter than every binary operator
        ExprP e = postfix();
        while (eat_op("|")) {
            if (peek().kind != Tok::NAME) throw TplError("expected a filter name");
            ExprP f = mk(Expr::FILTER);
            f->name = next().text;
            f->kids = {e};
            if (eat_op("(")) {
                while (!is_op(")")) {
                    f->kids.push_back(expr());
                    if (!eat_op(",")) break;
                }
                need_op(")");
            }
            e = f;
        }
        return e;
    }
    ExprP postfix() {
        ExprP e = primary();
        for (;;) {
            if (eat_op(".")) {
                if (peek().kind != Tok::NAME) throw TplError("expected a name after '.'");
                ExprP a = mk(Expr::ATTR); a->name = next().text; a->kids = {e};
                e = a;
            } else if (eat_op("[")) {
                ExprP lo, hi;
                bool slice = false;
                if (!is_op(":")) lo = expr();
                if (eat_op(":")) {
                    slice = true;
                    if (!is_op("]")) hi = expr();
                    if (is_op(":")) throw TplError("slice steps are not supported");
                }
                need_op("]");
                if (slice) { ExprP s = mk(Expr::SLICE); s->kids = {e, lo, hi}; e = s; }
                else { ExprP s = mk(Expr::INDEX); s->kids = {e, lo}; e = s; }
            } else if (eat_op("(")) {
                ExprP c = mk(Expr::CALL);
                c->kids = {e};
                while (!is_op(")")) {
                    if (peek().kind == Tok::NAME && toks_[pos_ + 1].kind == Tok::OP && toks_[pos_ + 1].text == "=") {
                        const std::string k = next().text;
                        pos_++;
                        c->kwargs.emplace_back(k, expr());
                    } else {
                        c->kids.push_back(expr());
                    }
                    if (!eat_op(",")) break;
                }
                need_op(")");
                e = c;
            } else {
                return e;
            }
        }
    }
    ExprP primary() {
        const Tok t = next();
        if (t.kind == Tok::INT) { ExprP e = mk(Expr::LIT); e->lit = Value::integer(t.ival); return e; }
        if (t.kind == Tok::STR) {
            std::string s = t.text;
            while (peek().kind == Tok::STR) s += next().text;  // adjacent literals concatenate
            ExprP e = mk(Expr::LIT); e->lit = Value::str(std::move(s)); return e;
        }
        if (t.kind == Tok::NAME) {
            ExprP e = mk(Expr::LIT);
            if (t.text == "true" || t.text == "True") { e->lit = Value::boolean(true); return e; }
            if (t.text == "false" || t.text == "False") { e->lit = Value::boolean(false); return e; }
            if (t.text == "none" || t.text == "None") { e->lit = Value::none(); return e; }
            e->kind = Expr::NAME; e->name = t.text;
            return e;
        }
        if (t.kind == Tok::OP && t.text == "(") { ExprP e = expr(); need_op(")"); return e; }
        if (t.kind == Tok::OP && t.text == "[") {
            ExprP l = mk(Expr::LISTLIT);
            while (!is_op("]")) { l->kids.push_back(expr()); if (!eat_op(",")) break; }
            need_op("]");
            return l;
        }
        if (t.kind == Tok::OP && t.text == "{") {
            ExprP d = mk(Expr::DICTLIT);
            while (!is_op("}")) {
                if (peek().kind != Tok::STR) throw TplError("mapping keys must be string literals");
                const std::string k = next().text;
                need_op(":");
                d->kwargs.emplace_back(k, expr());
                if (!eat_op(",")) break;
            }
            need_op("}");
            return d;
        }
        throw TplError("unexpected " + (t.kind == Tok::END ? std::string("end of expression") : "'" + t.text + "'"));
    }
    std::vector<Tok> toks_;
    size_t pos_ = 0;
};

// ---------------------------------------------------------------- template -> statement tree
struct Node;
using NodeP = std::shared_ptr<Node>;
struct Node {
    enum Kind { TEXT, OUTPUT, FOR, IF, SET } kind = TEXT;
    std::string text;                 // TEXT; FOR: loop variable; SET: target name
    std::string attr;                 // SET: `set ns.attr = ...`
    ExprP expr;                       // OUTPUT / FOR iterable / SET value
    std::vector<NodeP> body;          // FOR body
    std::vector<std::pair<ExprP, std::vector<NodeP>>> branches;  // IF: (condition or null for else, body)
};

struct Segment {
    enum Kind { TEXT, EXPR, STMT } kind;
    std::string src;
};

// splits the source into text / {{ }} / {% %} pieces and applies whitespace control: `-` markers, and the environment
// options chat templates are rendered with (trim_blocks: the newline after a block tag goes; lstrip_blocks: blanks
// before a block tag at the start of a line go)
std::vector<Segment> split_template(const std::string &src) {
    // keep_trailing_newline is off in that environment: one newline at the very end of the source is not part of it
    std::string t = src;
    if (!t.empty() && t.back() == '\n') {
        t.pop_back();
        if (!t.empty() && t.back() == '\r') t.pop_back();
    }
    std::vector<Segment> out;
    size_t i = 0;
    const size_t n = t.size();
    std::string text;
    bool strip_next_left = false, trim_newline = false;
    auto flush_text = [&](bool rstrip_all, bool lstrip_block) {
        if (trim_newline && !text.empty() && text[0] == '\n') text.erase(0, 1);
        else if (trim_newline && text.size() > 1 && text[0] == '\r' && text[1] == '\n') text.erase(0, 2);
        if (strip_next_left) text = strip(text, true, false);
        if (rstrip_all) text = strip(text, false, true);
        else if (lstrip_block) {  // blanks since the last newline (or the start of the template, or the last tag)
            size_t b = text.size();
            while (b > 0 && (text[b - 1] == ' ' || text[b - 1] == '\t')) b--;
            if (b == 0 ? (out.empty() || out.back().kind == Segment::STMT) : text[b - 1] == '\n') text.erase(b);
        }
        if (!text.empty()) out.push_back({Segment::TEXT, text});
        text.clear();
        strip_next_left = trim_newline = false;
    };
    while (i < n) {
        if (t[i] == '{' && i + 1 < n && (t[i + 1] == '{' || t[i + 1] == '%' || t[i + 1] == '#')) {
            const char kind = t[i + 1];
            const char *close = kind == '{' ? "}}" : kind == '%' ? "%}" : "#}";
            size_t j = i + 2;
            bool lminus = false, lplus = false;
            if (j < n && t[j] == '-') { lminus = true; j++; }
            else if (j < n && t[j] == '+' && kind != '{') { lplus = true; j++; }
            // find the closing marker (string literals may contain it)
            size_t e = j;
            char quote = 0;
            for (; e + 1 < n; e++) {
                if (quote) {
                    if (t[e] == '\\') e++;
                    else if (t[e] == quote) quote = 0;
                } else if (kind != '#' && (t[e] == '\'' || t[e] == '"')) {
                    quote = t[e];
                } else if (t[e] == close[0] && t[e + 1] == close[1]) {
                    break;
                }
            }
            if (e + 1 >= n) throw TplError("unterminated tag");
            size_t inner_end = e;
            bool rminus = false;
            if (inner_end > j && t[inner_end - 1] == '-') { rminus = true; inner_end--; }
            flush_text(lminus, kind != '{' && !lplus);
            if (kind == '{') out.push_back({Segment::EXPR, t.substr(j, inner_end - j)});
            else if (kind == '%') out.push_back({Segment::STMT, t.substr(j, inner_end - j)});
            i = e + 2;
            strip_next_left = rminus;
            trim_newline = !rminus && kind != '{';
            continue;
        }
        text += t[i++];
    }
    flush_text(false, false);
    return out;
}

class TreeBuilder {
   public:
    explicit TreeBuilder(std::vector<Segment> segs) : segs_(std::move(segs)) {}
    std::vector<NodeP> build() {
        std::vector<NodeP> body = block({});
        if (pos_ < segs_.size()) throw TplError("unexpected '" + segs_[pos_].src + "'");
        return body;
    }

   private:
    static std::string first_word(const std::string &src) {
        size_t a = 0;
        while (a < src.size() && is_space(src[a])) a++;
        size_t b = a;
        while (b < src.size() && (std::isalnum((unsigned char)src[b]) || src[b] == '_')) b++;
        return src.substr(a, b - a);
    }
    // parses until one of `stops` (a statement keyword) is next; leaves pos_ at that statement
    std::vector<NodeP> block(const std::vector<std::string> &stops) {
        std::vector<NodeP> out;
        while (pos_ < segs_.size()) {
            const Segment &sg = segs_[pos_];
            if (sg.kind == Segment::TEXT) {
                auto nd = std::make_shared<Node>(); nd->kind = Node::TEXT; nd->text = sg.src;
                out.push_back(nd); pos_++;
            } else if (sg.kind == Segment::EXPR) {
                auto nd = std::make_shared<Node>(); nd->kind = Node::OUTPUT; nd->expr = Parser(sg.src).parse_all();
                out.push_back(nd); pos_++;
            } else {
                const std::string w = first_word(sg.src);
                for (const std::string &s : stops)
                    if (w == s) return out;
                pos_++;
                if (w == "for") out.push_back(for_(sg.src));
                else if (w == "if") out.push_back(if_(sg.src));
                else if (w == "set") out.push_back(set_(sg.src));
                else if (w == "generation" || w == "endgeneration") continue;  // transformers' training-mask markers: no output
                else throw TplError("unsupported statement '" + w + "'");
            }
        }
        if (!stops.empty()) throw TplError("missing '" + stops.back() + "'");
        return out;
    }
    NodeP for_(const std::string &src) {
        Parser p(src);
        p.next();  // for
        if (p.peek().kind != Tok::NAME) throw TplError("expected a loop variable");
        auto nd = std::make_shared<Node>(); nd->kind = Node::FOR; nd->text = p.next().text;
        if (p.is_op(",")) throw TplError("tuple unpacking in for loops is not supported");
        if (!p.eat_name("in")) throw TplError("expected 'in'");
        nd->expr = p.expr();
        if (p.peek().kind != Tok::END) throw TplError("loop filters / recursive loops are not supported");
        nd->body = block({"endfor", "else"});
        if (first_word(segs_[pos_].src) == "else") throw TplError("for ... else is not supported");
        pos_++;
        return nd;
    }
    NodeP if_(const std::string &src) {
        auto nd = std::make_shared<Node>(); nd->kind = Node::IF;
        Parser p(src);
        p.next();  // if
        ExprP cond = p.expr();
        if (p.peek().kind != Tok::END) throw TplError("unexpected '" + p.peek().text + "' after the condition");
        for (;;) {
            std::vector<NodeP> body = block({"elif", "else", "endif"});
            nd->branches.emplace_back(cond, std::move(body));
            const std::string stmt = segs_[pos_].src, w = first_word(stmt);
            pos_++;
            if (w == "endif") break;
            if (w == "else") {
                std::vector<NodeP> eb = block({"endif"});
                nd->branches.emplace_back(nullptr, std::move(eb));
                pos_++;
                break;
            }
            Parser q(stmt);
            q.next();  // elif
            cond = q.expr();
            if (q.peek().kind != Tok::END) throw TplError("unexpected '" + q.peek().text + "' after the condition");
        }
        return nd;
    }
    NodeP set_(const std::string &src) {
        Parser p(src);
        p.next();  // set
        if (p.peek().kind != Tok::NAME) throw TplError("expected a name after 'set'");
        auto nd = std::make_shared<Node>(); nd->kind = Node::SET; nd->text = p.next().text;
        if (p.eat_op(".")) {
            if (p.peek().kind != Tok::NAME) throw TplError("expected an attribute name");
            nd->attr = p.next().text;
        }
        if (!p.eat_op("=")) throw TplError("block assignments are not supported");
        nd->expr = p.expr();
        if (p.peek().kind != Tok::END) throw TplError("unexpected '" + p.peek().text + "' in set");
        return nd;
    }
    std::vector<Segment> segs_;
    size_t pos_ = 0;
};

// ---------------------------------------------------------------- evaluation
class Interp {
   public:
    std::vector<std::map<std::string, Value>> scopes{1};
    std::string out;

    Value lookup(const std::string &name) const {
        for (size_t k = scopes.size(); k-- > 0;) {
            auto it = scopes[k].find(name);
            if (it != scopes[k].end()) return it->second;
        }
        return Value();
    }
    void run(const std::vector<NodeP> &body) {
        for (const NodeP &nd : body) {
            switch (nd->kind) {
                case Node::TEXT: out += nd->text; break;
                case Node::OUTPUT: out += eval(nd->expr).to_str(); break;
                case Node::SET: {
                    Value v = eval(nd->expr);
                    if (nd->attr.empty()) scopes.back()[nd->text] = std::move(v);
                    else {
                        Value ns = lookup(nd->text);
                        if (ns.kind != Value::DICT) throw TplError("cannot set an attribute of '" + nd->text + "'");
                        (*ns.dict)[nd->attr] = std::move(v);
                    }
                    break;
                }
                case Node::IF:
                    for (const auto &br : nd->branches)
                        if (!br.first || eval(br.first).truthy()) { run(br.second); break; }
                    break;
                case Node::FOR: {
                    const Value it = eval(nd->expr);
                    List items;
                    if (it.kind == Value::LIST) items = *it.list;
                    else if (it.kind == Value::STR) for (char c : it.s) items.push_back(Value::str(std::string(1, c)));
                    else if (it.kind == Value::DICT) for (const auto &kv : *it.dict) items.push_back(Value::str(kv.first));
                    else if (it.kind != Value::UNDEF && it.kind != Value::NONE) throw TplError("cannot iterate over this value");
                    const int64_t n = (int64_t)items.size();
                    for (int64_t k = 0; k < n; k++) {
                        scopes.emplace_back();
                        scopes.back()[nd->text] = items[(size_t)k];
                        Dict loop;
                        loop["index"] = Value::integer(k + 1); loop["index0"] = Value::integer(k);
                        loop["revindex"] = Value::integer(n - k); loop["revindex0"] = Value::integer(n - k - 1);
                        loop["first"] = Value::boolean(k == 0); loop["last"] = Value::boolean(k == n - 1);
                        loop["length"] = Value::integer(n);
                        if (k > 0) loop["previtem"] = items[(size_t)k - 1];
                        if (k + 1 < n) loop["nextitem"] = items[(size_t)k + 1];
                        scopes.back()["loop"] = Value::mkdict(std::move(loop));
                        run(nd->body);
                        scopes.pop_back();
                    }
                    break;
                }
            }
        }
    }

    Value eval(const ExprP &e) {
        switch (e->kind) {
            case Expr::LIT: return e->lit;
            case Expr::NAME: return lookup(e->name);
            case Expr::LISTLIT: {
                List l;
                for (const ExprP &k : e->kids) l.push_back(eval(k));
                return Value::mklist(std::move(l));
            }
            case Expr::DICTLIT: {
                Dict d;
                for (const auto &kv : e->kwargs) d[kv.first] = eval(kv.second);
                return Value::mkdict(std::move(d));
            }
            case Expr::ATTR: {
                const Value base = eval(e->kids[0]);
                if (base.kind == Value::DICT) {
                    auto it = base.dict->find(e->name);
                    return it == base.dict->end() ? Value() : it->second;
                }
                return Value();  // an attribute of anything else is undefined (methods are handled in CALL)
            }
            case Expr::INDEX: {
                const Value base = eval(e->kids[0]), idx = eval(e->kids[1]);
                if (base.kind == Value::DICT && idx.kind == Value::STR) {
                    auto it = base.dict->find(idx.s);
                    return it == base.dict->end() ? Value() : it->second;
                }
                if ((base.kind == Value::LIST || base.kind == Value::STR) && idx.kind == Value::INT) {
                    const int64_t n = base.kind == Value::LIST ? (int64_t)base.list->size() : (int64_t)base.s.size();
                    const int64_t k = idx.i < 0 ? idx.i + n : idx.i;
                    if (k < 0 || k >= n) return Value();
                    return base.kind == Value::LIST ? (*base.list)[(size_t)k] : Value::str(std::string(1, base.s[(size_t)k]));
                }
                return Value();
            }
            case Expr::SLICE: {
                const Value base = eval(e->kids[0]);
                if (base.kind != Value::LIST && base.kind != Value::STR) throw TplError("cannot slice this value");
                const int64_t n = base.kind == Value::LIST ? (int64_t)base.list->size() : (int64_t)base.s.size();
                auto bound = [&](const ExprP &x, int64_t dflt) {
                    if (!x) return dflt;
                    const Value v = eval(x);
                    if (v.kind == Value::NONE || v.kind == Value::UNDEF) return dflt;
                    if (v.kind != Value::INT) throw TplError("slice bounds must be integers");
                    int64_t k = v.i < 0 ? v.i + n : v.i;
                    return k < 0 ? (int64_t)0 : k > n ? n : k;
                };
                const int64_t lo = bound(e->kids[1], 0), hi = std::max(lo, bound(e->kids[2], n));
                if (base.kind == Value::STR) return Value::str(base.s.substr((size_t)lo, (size_t)(hi - lo)));
                return Value::mklist(List(base.list->begin() + lo, base.list->begin() + hi));
            }
            case Expr::UNARY: {
                const Value v = eval(e->kids[0]);
                if (e->name == "not") return Value::boolean(!v.truthy());
                if (v.kind != Value::INT) throw TplError("unary " + e->name + " needs an integer");
                return Value::integer(e->name == "-" ? -v.i : v.i);
            }
            case Expr::COND: return eval(e->kids[0]).truthy() ? eval(e->kids[1]) : eval(e->kids[2]);
            case Expr::BINARY: return binary(e);
            case Expr::TEST: {
                const Value v = eval(e->kids[0]);
                bool r;
                if (e->name == "defined") r = v.kind != Value::UNDEF;
                else if (e->name == "undefined") r = v.kind == Value::UNDEF;
                else if (e->name == "none") r = v.kind == Value::NONE;
                else if (e->name == "string") r = v.kind == Value::STR;
                else if (e->name == "mapping") r = v.kind == Value::DICT;
                else if (e->name == "iterable") r = v.kind == Value::LIST || v.kind == Value::STR || v.kind == Value::DICT;
                else if (e->name == "sequence") r = v.kind == Value::LIST || v.kind == Value::STR;
                else if (e->name == "number" || e->name == "integer") r = v.kind == Value::INT;
                else if (e->name == "boolean") r = v.kind == Value::BOOL;
                else if (e->name == "true") r = v.kind == Value::BOOL && v.b;
                else if (e->name == "false") r = v.kind == Value::BOOL && !v.b;
                else throw TplError("unsupported test '" + e->name + "'");
                return Value::boolean(r != e->negate);
            }
            case Expr::FILTER: return filter(e);
            case Expr::CALL: return call(e);
        }
        throw TplError("bad expression");
    }

   private:
    Value binary(const ExprP &e) {
        const std::string &op = e->name;
        if (op == "and") { const Value a = eval(e->kids[0]); return a.truthy() ? eval(e->kids[1]) : a; }
        if (op == "or") { const Value a = eval(e->kids[0]); return a.truthy() ? a : eval(e->kids[1]); }
        const Value a = eval(e->kids[0]), b = eval(e->kids[1]);
        if (op == "==") return Value::boolean(equal(a, b));
        if (op == "!=") return Value::boolean(!equal(a, b));
        if (op == "~") return Value::str(a.to_str() + b.to_str());
        if (op == "in") {
            if (b.kind == Value::STR && a.kind == Value::STR) return Value::boolean(b.s.find(a.s) != std::string::npos);
            if (b.kind == Value::LIST) {
                for (const Value &x : *b.list)
                    if (equal(a, x)) return Value::boolean(true);
                return Value::boolean(false);
            }
            if (b.kind == Value::DICT && a.kind == Value::STR) return Value::boolean(b.dict->count(a.s) != 0);
            throw TplError("'in' needs a string, a list or a mapping on the right");
        }
        if (op == "+") {
            if (a.kind == Value::STR && b.kind == Value::STR) return Value::str(a.s + b.s);
            if (a.kind == Value::LIST && b.kind == Value::LIST) { List l = *a.list; l.insert(l.end(), b.list->begin(), b.list->end()); return Value::mklist(std::move(l)); }
        }
        if (op == "<" || op == ">" || op == "<=" || op == ">=") {
            int c;
            if (a.kind == Value::INT && b.kind == Value::INT) c = a.i < b.i ? -1 : a.i > b.i ? 1 : 0;
            else if (a.kind == Value::STR && b.kind == Value::STR) c = a.s.compare(b.s);
            else throw TplError("cannot order these values");
            return Value::boolean(op == "<" ? c < 0 : op == ">" ? c > 0 : op == "<=" ? c <= 0 : c >= 0);
        }
        if (a.kind == Value::INT && b.kind == Value::INT) {
            if (op == "+") return Value::integer(a.i + b.i);
            if (op == "-") return Value::integer(a.i - b.i);
            if (op == "*") return Value::integer(a.i * b.i);
            if (op == "//" || op == "%") {
                if (b.i == 0) throw TplError("division by zero");
                int64_t q = a.i / b.i, r = a.i % b.i;
                if (r != 0 && ((r < 0) != (b.i < 0))) { q--; r += b.i; }  // Python's floor semantics
                return Value::integer(op == "%" ? r : q);
            }
        }
        if (op == "*" && a.kind == Value::STR && b.kind == Value::INT) {
            std::string s;
            for (int64_t k = 0; k < b.i; k++) s += a.s;
            return Value::str(std::move(s));
        }
        throw TplError("unsupported operands for '" + op + "'");
    }
    Value filter(const ExprP &e) {
        const Value v = eval(e->kids[0]);
        const std::string &f = e->name;
        if (f == "trim") return Value::str(strip(v.to_str(), true, true));
        if (f == "length" || f == "count") {
            if (v.kind == Value::STR) return Value::integer((int64_t)v.s.size());
            if (v.kind == Value::LIST) return Value::integer((int64_t)v.list->size());
            if (v.kind == Value::DICT) return Value::integer((int64_t)v.dict->size());
            throw TplError("length of a value without one");
        }
        if (f == "upper" || f == "lower") {
            std::string s = v.to_str();
            for (char &c : s) c = f == "upper" ? (char)std::toupper((unsigned char)c) : (char)std::tolower((unsigned char)c);
            return Value::str(std::move(s));
        }
        if (f == "string") return Value::str(v.to_str());
        if (f == "safe") return v;
        if (f == "first" || f == "last") {
            if (v.kind == Value::LIST) return v.list->empty() ? Value() : f == "first" ? v.list->front() : v.list->back();
            if (v.kind == Value::STR) return v.s.empty() ? Value() : Value::str(std::string(1, f == "first" ? v.s.front() : v.s.back()));
            throw TplError(f + " of a value that is no sequence");
        }
        if (f == "default" || f == "d") {
            const Value dflt = e->kids.size() > 1 ? eval(e->kids[1]) : Value::str("");
            const bool falsy_too = e->kids.size() > 2 && eval(e->kids[2]).truthy();
            return v.kind == Value::UNDEF || (falsy_too && !v.truthy()) ? dflt : v;
        }
        if (f == "join") {
            if (v.kind != Value::LIST) throw TplError("join needs a list");
            const std::string sep = e->kids.size() > 1 ? eval(e->kids[1]).to_str() : "";
            std::string s;
            for (size_t k = 0; k < v.list->size(); k++) s += (k ? sep : "") + (*v.list)[k].to_str();
            return Value::str(std::move(s));
        }
        if (f == "int") {
            if (v.kind == Value::INT) return v;
            if (v.kind == Value::BOOL) return Value::integer(v.b ? 1 : 0);
            try { return Value::integer(std::stoll(v.to_str())); } catch (...) { return Value::integer(0); }
        }
        throw TplError("unsupported filter '" + f + "'");
    }
    Value call(const ExprP &e) {
        const ExprP &fn = e->kids[0];
        std::vector<Value> args;
        for (size_t k = 1; k < e->kids.size(); k++) args.push_back(eval(e->kids[k]));
        if (fn->kind == Expr::NAME) {
            if (fn->name == "raise_exception") throw TplError("the template raised: " + (args.empty() ? std::string() : args[0].to_str()));
            if (fn->name == "namespace") {
                Dict d;
                for (const auto &kv : e->kwargs) d[kv.first] = eval(kv.second);
                return Value::mkdict(std::move(d));
            }
            if (fn->name == "range") {
                int64_t lo = 0, hi = 0;
                if (args.size() == 1 && args[0].kind == Value::INT) hi = args[0].i;
                else if (args.size() == 2 && args[0].kind == Value::INT && args[1].kind == Value::INT) { lo = args[0].i; hi = args[1].i; }
                else throw TplError("range takes one or two integers");
                List l;
                for (int64_t k = lo; k < hi; k++) l.push_back(Value::integer(k));
                return Value::mklist(std::move(l));
            }
            throw TplError("unsupported function '" + fn->name + "'");
        }
        if (fn->kind == Expr::ATTR) {  // a method of a string / mapping
            const Value self = eval(fn->kids[0]);
            const std::string &m = fn->name;
            if (self.kind == Value::STR) {
                if (m == "strip" && args.empty()) return Value::str(strip(self.s, true, true));
                if (m == "lstrip" && args.empty()) return Value::str(strip(self.s, true, false));
                if (m == "rstrip" && args.empty()) return Value::str(strip(self.s, false, true));
                if (m == "upper" || m == "lower") {
                    std::string s = self.s;
                    for (char &c : s) c = m == "upper" ? (char)std::toupper((unsigned char)c) : (char)std::tolower((unsigned char)c);
                    return Value::str(std::move(s));
                }
                if ((m == "startswith" || m == "endswith") && args.size() == 1 && args[0].kind == Value::STR) {
                    const std::string &p = args[0].s;
                    if (p.size() > self.s.size()) return Value::boolean(false);
                    return Value::boolean(m == "startswith" ? self.s.compare(0, p.size(), p) == 0
                                                            : self.s.compare(self.s.size() - p.size(), p.size(), p) == 0);
                }
                if (m == "split" && args.size() <= 1) {
                    List l;
                    if (args.empty()) {
                        size_t a = 0;
                        while (a < self.s.size()) {
                            while (a < self.s.size() && is_space(self.s[a])) a++;
                            size_t b = a;
                            while (b < self.s.size() && !is_space(self.s[b])) b++;
                            if (b > a) l.push_back(Value::str(self.s.substr(a, b - a)));
                            a = b;
                        }
                    } else {
                        const std::string sep = args[0].to_str();
                        if (sep.empty()) throw TplError("empty separator");
                        size_t a = 0;
                        for (;;) {
                            const size_t b = self.s.find(sep, a);
                            l.push_back(Value::str(self.s.substr(a, b == std::string::npos ? std::string::npos : b - a)));
                            if (b == std::string::npos) break;
                            a = b + sep.size();
                        }
                    }
                    return Value::mklist(std::move(l));
                }
            }
            if (self.kind == Value::DICT && m == "get" && !args.empty() && args[0].kind == Value::STR) {
                auto it = self.dict->find(args[0].s);
                if (it != self.dict->end()) return it->second;
                return args.size() > 1 ? args[1] : Value::none();
            }
            throw TplError("unsupported method '" + m + "'");
        }
        throw TplError("unsupported call");
    }
};

}  // namespace

bool render_chat_template(const std::string &tmpl, const std::vector<ChatMessage> &messages, const std::string &bos_token,
                          const std::string &eos_token, bool add_generation_prompt, std::string *out, std::string *err) {
    try {
        const std::vector<NodeP> tree = TreeBuilder(split_template(tmpl)).build();
        Interp in;
        List msgs;
        for (const ChatMessage &m : messages) {
            Dict d;
            d["role"] = Value::str(m.role);
            d["content"] = Value::str(m.content);
            msgs.push_back(Value::mkdict(std::move(d)));
        }
        in.scopes[0]["messages"] = Value::mklist(std::move(msgs));
        in.scopes[0]["bos_token"] = Value::str(bos_token);
        in.scopes[0]["eos_token"] = Value::str(eos_token);
        in.scopes[0]["add_generation_prompt"] = Value::boolean(add_generation_prompt);
        in.run(tree);
        *out = std::move(in.out);
        return true;
    } catch (const std::exception &e) {
        if (err) *err = e.what();
        return false;
    }
}

}  // namespace aios
