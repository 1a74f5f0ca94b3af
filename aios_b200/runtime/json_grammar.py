"""Incremental JSON-object acceptor for constrained sampling (``response_format: {"type": "json_object"}``).

The reference's unary ``Infer`` always asks llama-server for JSON mode (runtime/src/inference.rs:119-121) and the
orchestrator parses the answer as JSON.  llama-server implements it as a grammar
(``root ::= object``) that masks the sampler's candidates; this module is that grammar as a byte-level pushdown
automaton: ``JsonPrefix.feed(bytes)`` answers whether the text so far is still a prefix of one JSON object, and
``complete`` says when the object has closed.  The sampler driver (inference.py) keeps only the candidates whose
piece the automaton accepts.

Grammar (RFC 8259, as llama.cpp's json.gbnf restricts it): root = ws object ws; value = object | array | string |
number | true | false | null; strings with the standard escapes and \\uXXXX; numbers -?(0|[1-9][0-9]*)(.[0-9]+)?
([eE][+-]?[0-9]+)?; whitespace = space, tab, CR, LF.  Bytes >= 0x80 are accepted inside strings when they form
well-formed UTF-8 sequences.
"""
from __future__ import annotations

_WS = b" \t\r\n"
_DIGITS = b"0123456789"
_HEX = b"0123456789abcdefABCDEF"

# modes
(_VALUE, _OBJ_FIRST_KEY, _OBJ_KEY, _OBJ_COLON, _OBJ_NEXT, _ARR_FIRST, _ARR_NEXT, _STRING, _STR_ESC, _STR_HEX,
 _NUMBER, _LITERAL, _DONE, _START) = range(14)

# number sub-states
(_N_MINUS, _N_ZERO, _N_INT, _N_DOT, _N_FRAC, _N_E, _N_ESIGN, _N_EXP) = range(8)


class JsonPrefix:
    """State of the automaton; cheap to copy (``clone``) so candidates can be tried and discarded."""

    __slots__ = ("stack", "mode", "sub", "lit", "key_string", "utf8_left", "depth_limit")

    def __init__(self, depth_limit: int = 64):
        self.stack: list[int] = []      # 0 = inside an object, 1 = inside an array
        self.mode = _START
        self.sub = 0                    # hex digits left / number sub-state / literal position
        self.lit = b""
        self.key_string = False         # the string being read is an object key
        self.utf8_left = 0              # continuation bytes still expected inside a string
        self.depth_limit = depth_limit

    def clone(self) -> "JsonPrefix":
        c = JsonPrefix.__new__(JsonPrefix)
        c.stack = list(self.stack)
        c.mode, c.sub, c.lit, c.key_string = self.mode, self.sub, self.lit, self.key_string
        c.utf8_left, c.depth_limit = self.utf8_left, self.depth_limit
        return c

    @property
    def complete(self) -> bool:
        """The top-level object has closed (only whitespace may follow)."""
        return self.mode == _DONE

    # ------------------------------------------------------------------ one byte
    def _after_value(self):
        if not self.stack:
            self.mode = _DONE
        else:
            self.mode = _OBJ_NEXT if self.stack[-1] == 0 else _ARR_NEXT

    def _start_value(self, b: int) -> bool:
        if b == 0x7B:  # {
            if len(self.stack) >= self.depth_limit:
                return False
            self.stack.append(0)
            self.mode = _OBJ_FIRST_KEY
        elif b == 0x5B:  # [
            if len(self.stack) >= self.depth_limit:
                return False
            self.stack.append(1)
            self.mode = _ARR_FIRST
        elif b == 0x22:  # "
            self.mode, self.key_string = _STRING, False
        elif b == 0x2D:  # -
            self.mode, self.sub = _NUMBER, _N_MINUS
        elif b == 0x30:
            self.mode, self.sub = _NUMBER, _N_ZERO
        elif 0x31 <= b <= 0x39:
            self.mode, self.sub = _NUMBER, _N_INT
        elif b == 0x74:
            self.mode, self.lit, self.sub = _LITERAL, b"true", 1
        elif b == 0x66:
            self.mode, self.lit, self.sub = _LITERAL, b"false", 1
        elif b == 0x6E:
            self.mode, self.lit, self.sub = _LITERAL, b"null", 1
        else:
            return False
        return True

    def _number_can_end(self) -> bool:
        return self.sub in (_N_ZERO, _N_INT, _N_FRAC, _N_EXP)

    def _byte(self, b: int) -> bool:
        m = self.mode
        if m == _STRING:
            if self.utf8_left:
                if 0x80 <= b <= 0xBF:
                    self.utf8_left -= 1
                    return True
                return False
            if b == 0x22:
                if self.key_string:
                    self.mode = _OBJ_COLON
                else:
                    self._after_value()
                return True
            if b == 0x5C:
                self.mode = _STR_ESC
                return True
            if b < 0x20:
                return False
            if b < 0x80:
                return True
            if 0xC2 <= b <= 0xDF:
                self.utf8_left = 1
            elif 0xE0 <= b <= 0xEF:
                self.utf8_left = 2
            elif 0xF0 <= b <= 0xF4:
                self.utf8_left = 3
            else:
                return False
            return True
        if m == _STR_ESC:
            if b in b'"\\/bfnrt':
                self.mode = _STRING
                return True
            if b == 0x75:  # u
                self.mode, self.sub = _STR_HEX, 4
                return True
            return False
        if m == _STR_HEX:
            if b in _HEX:
                self.sub -= 1
                if self.sub == 0:
                    self.mode = _STRING
                return True
            return False
        if m == _NUMBER:
            s = self.sub
            if s == _N_MINUS:
                if b == 0x30:
                    self.sub = _N_ZERO
                    return True
                if 0x31 <= b <= 0x39:
                    self.sub = _N_INT
                    return True
                return False
            if s in (_N_INT, _N_FRAC, _N_EXP) and b in _DIGITS:
                return True
            if s in (_N_ZERO, _N_INT) and b == 0x2E:
                self.sub = _N_DOT
                return True
            if s in (_N_ZERO, _N_INT, _N_FRAC) and b in b"eE":
                self.sub = _N_E
                return True
            if s == _N_DOT:
                if b in _DIGITS:
                    self.sub = _N_FRAC
                    return True
                return False
            if s == _N_E:
                if b in b"+-":
                    self.sub = _N_ESIGN
                    return True
                if b in _DIGITS:
                    self.sub = _N_EXP
                    return True
                return False
            if s == _N_ESIGN:
                if b in _DIGITS:
                    self.sub = _N_EXP
                    return True
                return False
            # anything else ends the number, if it may end here, and is then looked at again
            if not self._number_can_end():
                return False
            self._after_value()
            return self._byte(b)
        if m == _LITERAL:
            if b == self.lit[self.sub]:
                self.sub += 1
                if self.sub == len(self.lit):
                    self._after_value()
                return True
            return False
        # structural modes: whitespace is free
        if b in _WS:
            return True
        if m == _START:
            if b == 0x7B:
                self.stack.append(0)
                self.mode = _OBJ_FIRST_KEY
                return True
            return False
        if m == _VALUE:
            return self._start_value(b)
        if m == _OBJ_FIRST_KEY or m == _OBJ_KEY:
            if b == 0x22:
                self.mode, self.key_string = _STRING, True
                return True
            if b == 0x7D and m == _OBJ_FIRST_KEY:
                self.stack.pop()
                self._after_value()
                return True
            return False
        if m == _OBJ_COLON:
            if b == 0x3A:
                self.mode = _VALUE
                return True
            return False
        if m == _OBJ_NEXT:
            if b == 0x2C:
                self.mode = _OBJ_KEY
                return True
            if b == 0x7D:
                self.stack.pop()
                self._after_value()
                return True
            return False
        if m == _ARR_FIRST:
            if b == 0x5D:
                self.stack.pop()
                self._after_value()
                return True
            return self._start_value(b)
        if m == _ARR_NEXT:
            if b == 0x2C:
                self.mode = _VALUE
                return True
            if b == 0x5D:
                self.stack.pop()
                self._after_value()
                return True
            return False
        return False  # _DONE: only whitespace (handled above)

    # ------------------------------------------------------------------ public
    def feed(self, data: bytes) -> bool:
        """Advance over ``data``; False (state then undefined — work on a clone) if it cannot extend a JSON object."""
        for b in data:
            if not self._byte(b):
                return False
        return True

    def accepts(self, data: bytes) -> bool:
        """Would ``data`` be accepted?  The state is left untouched."""
        return bool(data) and self.clone().feed(data)


def is_json_object_prefix(text: bytes) -> bool:
    return JsonPrefix().feed(text)
