"""ctypes binding of libpcre2-8 (10.42 in this image): a third, independent backtracking engine for
pinning the oracle's java.util.regex restatement. Test infrastructure only.

PCRE2 is the closest dialect to java.util.regex among the engines available here: possessive
quantifiers, atomic groups, variable-alternative look-behind, `\\Q..\\E`, `\\k<name>`, `\\x{h..h}`,
`\\e`, `\\cX`, `\\h \\v \\R` all parse natively, which CPython `re`/`regex` cannot do. Compiled with
PCRE2_UTF and PCRE2_NEWLINE_ANY so that `.` and `$` use Unicode line terminators as Java does; the two
differences that remain are kept out of the texts: PCRE2's "any newline" also counts VT and FF, and
its caseless matching of non-ASCII characters folds by Unicode where Java without (?u) does not.
PCRE2's own optimisations (auto-possessification, start-of-match shortcuts) are switched off: 10.42
returned a wrong 'no match' with them on one generated pattern.
"""
from __future__ import annotations

import ctypes as C
import ctypes.util

PCRE2_CASELESS = 0x00000008
PCRE2_UTF = 0x00080000
PCRE2_NO_AUTO_POSSESS = 0x00004000
PCRE2_NO_DOTSTAR_ANCHOR = 0x00008000
PCRE2_NO_START_OPTIMIZE = 0x00010000
PCRE2_NEWLINE_ANY = 4
PCRE2_ERROR_NOMATCH = -1


def _load():
    for name in ("libpcre2-8.so.0", ctypes.util.find_library("pcre2-8")):
        if not name:
            continue
        try:
            return C.CDLL(name)
        except OSError:
            continue
    return None


_L = _load()
available = _L is not None

if available:
    _L.pcre2_compile_8.restype = C.c_void_p
    _L.pcre2_compile_8.argtypes = [C.c_char_p, C.c_size_t, C.c_uint32, C.POINTER(C.c_int), C.POINTER(C.c_size_t), C.c_void_p]
    _L.pcre2_compile_context_create_8.restype = C.c_void_p
    _L.pcre2_compile_context_create_8.argtypes = [C.c_void_p]
    _L.pcre2_set_newline_8.restype = C.c_int
    _L.pcre2_set_newline_8.argtypes = [C.c_void_p, C.c_uint32]
    _L.pcre2_match_data_create_from_pattern_8.restype = C.c_void_p
    _L.pcre2_match_data_create_from_pattern_8.argtypes = [C.c_void_p, C.c_void_p]
    _L.pcre2_match_8.restype = C.c_int
    _L.pcre2_match_8.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t, C.c_size_t, C.c_uint32, C.c_void_p, C.c_void_p]
    _L.pcre2_get_ovector_pointer_8.restype = C.POINTER(C.c_size_t)
    _L.pcre2_get_ovector_pointer_8.argtypes = [C.c_void_p]
    _L.pcre2_code_free_8.argtypes = [C.c_void_p]
    _L.pcre2_match_data_free_8.argtypes = [C.c_void_p]
    _L.pcre2_get_error_message_8.argtypes = [C.c_int, C.c_char_p, C.c_size_t]
    _L.pcre2_config_8.argtypes = [C.c_uint32, C.c_void_p]
    _L.pcre2_pattern_info_8.restype = C.c_int
    _L.pcre2_pattern_info_8.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
    _CCTX = _L.pcre2_compile_context_create_8(None)
    _L.pcre2_set_newline_8(_CCTX, PCRE2_NEWLINE_ANY)


class Pcre2Error(ValueError):
    pass


def version() -> str:
    buf = C.create_string_buffer(64)
    _L.pcre2_config_8(11, buf)  # PCRE2_CONFIG_VERSION
    return buf.value.decode()


class Regex:
    """One compiled pattern; `search(text)` is Matcher.find() from position 0: None or the (start, end)
    span in code points."""

    def __init__(self, pattern: str, caseless: bool = False):
        pat = pattern.encode("utf-8")
        err, off = C.c_int(0), C.c_size_t(0)
        self.code = _L.pcre2_compile_8(pat, len(pat), PCRE2_UTF | PCRE2_NO_AUTO_POSSESS | PCRE2_NO_DOTSTAR_ANCHOR | PCRE2_NO_START_OPTIMIZE | (PCRE2_CASELESS if caseless else 0), C.byref(err), C.byref(off), _CCTX)
        if not self.code:
            buf = C.create_string_buffer(256)
            _L.pcre2_get_error_message_8(err.value, buf, 256)
            raise Pcre2Error(f"{buf.value.decode()} at {off.value}")
        self.md = _L.pcre2_match_data_create_from_pattern_8(self.code, None)

    def search(self, text: str):
        b = text.encode("utf-8")
        rc = _L.pcre2_match_8(self.code, b, len(b), 0, 0, self.md, None)
        if rc == PCRE2_ERROR_NOMATCH:
            return None
        if rc < 0:      # -47 match limit, -53 depth limit, ...: the caller skips the case
            raise Pcre2Error(f"pcre2_match error {rc}")
        ov = _L.pcre2_get_ovector_pointer_8(self.md)
        s, e = ov[0], ov[1]
        return len(b[:s].decode("utf-8")), len(b[:e].decode("utf-8"))

    def can_match_empty(self) -> bool:
        v = C.c_uint32(0)
        _L.pcre2_pattern_info_8(self.code, 13, C.byref(v))  # PCRE2_INFO_MATCHEMPTY
        return bool(v.value)

    def __del__(self):
        try:
            if getattr(self, "md", None):
                _L.pcre2_match_data_free_8(self.md)
            if getattr(self, "code", None):
                _L.pcre2_code_free_8(self.code)
        except Exception:
            pass
