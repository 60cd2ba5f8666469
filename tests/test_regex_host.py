"""CPU tests of AddKeywordPattern's regex subset (preprocessor.go:176-184 + :146): the product's compiler and the host
build of its device walk (ordered-thread simulation) against two independent restatements of Go's leftmost-first,
non-overlapping FindAllString count - the C oracle's backtracking matcher and Python's `re` fed with explicit classes."""
import numpy as np
import pytest

from oracle import c_oracle as co
from oracle import py_oracle as po
from tests import hostcheck as hc
from tests import randregex

CASES = [
    (b"ur+gent", b"urgent urrrgent ugent urgen", 2),
    (b"colou?r", b"color colour colouur", 2),
    (b"(?i)a[bc]+d", b"ABCD abd ad aBcBd", 3),
    (b"abcd|ab", b"abcx abcd ab", 3),                  # leftmost-first: the left alternative wins where it matches
    (b"(a|ab)(c|bcd)", b"abcd abc ac", 3),
    (b"a+", b"aaa a aa", 3), (b"a+?", b"aaa", 3),        # greedy vs lazy changes the non-overlapping count
    (b"[0-9]{2,3}", b"1 12 123 1234 12345", 5),
    (b".", "é\n".encode() + b"\xff a", 4),         # one rune per ".", never "\n"; an invalid byte is one rune
    (b"(?s).", b"a\nb", 3),
    (b"[^a]", b"aba\n", 2),
    (b"\\d+\\.\\d+", b"3.14 2. .5 10.01", 2),
    (b"(?i)[r-t]+k", "RSTK stK ſK sk".encode(), 4),   # fold orbits of s and k inside a class
    (b"\\w+@\\w+\\.com", b"mail bob@x.com or a_b@yy.com!", 2),
    (b"x{3}", b"xxxxxxx", 2), (b"(?:ab)+c", b"ababc abc ab c", 2),
]


@pytest.mark.parametrize("pat,text,want", CASES)
def test_known_counts_three_ways(pat, text, want):
    assert hc.regex_count(pat, text) == want
    assert co.regex_count(pat, text) == want
    assert sum(1 for _ in po.go_regex_to_py(pat.decode()).finditer(po.decode(text))) == want


@pytest.mark.parametrize("bad", [b"a*", b"^a", b"a$", b"(a?)*b", b"\\bfoo", b"(?P<n>a)", b"[[:alpha:]]", b"a**", b"(", b"a)",
                                 "(?i)é".encode(), b"x{200}", b"(a|b|c|d|e|f){20}", b"", b"a|", b"(?m)a", b"\\1", b"[b-a]"])
def test_outside_the_subset_is_an_error_not_an_approximation(bad):
    with pytest.raises(ValueError):
        hc.regex_count(bad, b"aaa")


def test_random_patterns_three_ways():
    rng = np.random.default_rng(20261107)
    accepted = 0
    for _ in range(1500):
        pat = randregex.pattern(rng).encode()
        text = randregex.text(rng, int(rng.integers(1, 120)))
        try:
            got = hc.regex_count(pat, text)
        except ValueError:
            continue                      # nullable, too large, ...: refused, never approximated
        accepted += 1
        want_c = co.regex_count(pat, text)
        assert want_c is not None and got == want_c, (pat, text, got, want_c)
        try:
            rx = po.go_regex_to_py(pat.decode())
        except ValueError:
            continue
        want_py = sum(1 for _ in rx.finditer(po.decode(text)))
        assert got == want_py, (pat, text, got, want_py)
    assert accepted > 700
