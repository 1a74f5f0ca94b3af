// jinja_lite.h — the subset of Jinja2 that chat templates in GGUF files use (tokenizer.chat_template), host side, C++17.
//
// The reference sends a {system, user} message list (runtime/src/inference.rs:363-376) and llama-server renders it with
// the model's own template (its Jinja engine, or the built-in formats it recognises by marker strings).  This is the
// first route: the template text from the file is parsed and evaluated — statements for / if / elif / else / set,
// expressions with string, integer, boolean and none literals, names, attribute and subscript access, slices, calls of
// raise_exception and of the string methods strip / lstrip / rstrip / upper / lower / startswith / endswith, the
// filters trim / length / upper / lower / first / last / string / default, the tests defined / none / string, the
// operators + ~ - * % == != < > <= >= and or not in, the loop variable (index, index0, first, last, length), whitespace
// control ({%- -%} {{- -}}) and the environment options transformers (and llama.cpp after it) render with
// (trim_blocks, lstrip_blocks).  Anything else makes render_chat_template return false and the caller falls back to
// the marker-string formats (tokenizer.cpp).  Checked against the jinja2 package's output
// (tests/golden/chat_templates.json, scripts/make_golden_chat.py).
#pragma once
#include <string>
#include <vector>

namespace aios {

struct ChatMessage {
    std::string role, content;
};

// Renders `tmpl` with messages / bos_token / eos_token / add_generation_prompt defined.  false: *err says why (an
// unsupported construct, a syntax error, or the template's own raise_exception).
bool render_chat_template(const std::string &tmpl, const std::vector<ChatMessage> &messages, const std::string &bos_token,
                          const std::string &eos_token, bool add_generation_prompt, std::string *out, std::string *err);

}  // namespace aios
