"""Structural lint of the Java plug-in sources (no JDK on the build or GPU boxes, so javac never sees them here).

A small Java lexer (comments, string / char / text-block literals) feeds checks that catch what an editor without a
compiler lets through: unbalanced delimiters, a file whose public type or package does not match its path, imports of this
project's classes that do not exist in the flavour being built, statements that lost their semicolon before a closing
brace, and duplicate method declarations inside one class body.
"""
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent / "java"
FILES = sorted(ROOT.rglob("*.java"))


def lex(src: str):
    """Tokens [(kind, text, line)], kinds: id, num, str, chr, op. Raises ValueError on an unterminated literal/comment."""
    out, i, line, n = [], 0, 1, len(src)
    while i < n:
        c = src[i]
        if c == "\n":
            line += 1; i += 1
        elif c.isspace():
            i += 1
        elif src.startswith("//", i):
            j = src.find("\n", i)
            i = n if j < 0 else j
        elif src.startswith("/*", i):
            j = src.find("*/", i + 2)
            if j < 0:
                raise ValueError(f"unterminated comment at line {line}")
            line += src.count("\n", i, j); i = j + 2
        elif src.startswith('"""', i):
            j = src.find('"""', i + 3)
            if j < 0:
                raise ValueError(f"unterminated text block at line {line}")
            out.append(("str", src[i:j + 3], line)); line += src.count("\n", i, j); i = j + 3
        elif c in "\"'":
            j = i + 1
            while j < n and src[j] != c:
                if src[j] == "\n":
                    raise ValueError(f"unterminated literal at line {line}")
                j += 2 if src[j] == "\\" else 1
            if j >= n:
                raise ValueError(f"unterminated literal at line {line}")
            out.append(("str" if c == '"' else "chr", src[i:j + 1], line)); i = j + 1
        elif c.isalpha() or c in "_$@":
            j = i + 1
            while j < n and (src[j].isalnum() or src[j] in "_$"):
                j += 1
            out.append(("id", src[i:j], line)); i = j
        elif c.isdigit():
            j = i + 1
            while j < n and (src[j].isalnum() or src[j] in "._"):
                j += 1
            out.append(("num", src[i:j], line)); i = j
        else:
            out.append(("op", c, line)); i += 1
    return out


def test_the_lexer_itself():
    toks = lex('class A { String s = "a}\\"{"; char c = \'}\'; /* { */ int x = 1; // }\n}')
    assert [t[1] for t in toks if t[0] == "op" and t[1] in "{}"] == ["{", "}"]
    with pytest.raises(ValueError):
        lex('String s = "abc;\n')
    with pytest.raises(ValueError):
        lex("/* never closed")


@pytest.mark.parametrize("path", FILES, ids=lambda p: str(p.relative_to(ROOT)))
def test_delimiters_balance_and_statements_end(path):
    _check_delimiters(lex(path.read_text()), path)


def test_the_delimiter_check_catches_what_it_is_for():
    class Name:
        name = "Synthetic.java"
    good = "class A { int[] a = {f(1), g(2)}; void m() { if (x) { y(); } } }"
    _check_delimiters(lex(good), Name)
    for bad in ("class A { void m() { y() } }",            # statement lost its semicolon
                "class A { void m() { y(; } }",             # unbalanced parenthesis
                "class A { void m() { y(); }",              # unclosed brace
                "class A { void m() { y(); } } }"):         # one brace too many
        with pytest.raises(AssertionError):
            _check_delimiters(lex(bad), Name)


def _check_delimiters(toks, path):
    stack = []
    pairs = {")": "(", "]": "[", "}": "{"}
    for k, (kind, text, line) in enumerate(toks):
        if kind != "op":
            continue
        if text in "([{":
            stack.append((text, line))
        elif text in ")]}":
            assert stack and stack[-1][0] == pairs[text], f"{path.name}:{line}: unmatched {text!r}"
            stack.pop()
            if text == "}" and k > 0:
                prev = toks[k - 1]
                # before a closing brace comes ';' (statement), '}' (block), '{' (empty body), ',' (array / enum list) or
                # an enum constant / array element; an identifier or ')' here is a statement that lost its ';' --
                # except inside array initialisers and enum bodies, which are recognised by their opener
                if prev[0] == "op" and prev[1] == ")":
                    opener_is_init = _opener_is_initialiser(toks, k)
                    assert opener_is_init, f"{path.name}:{line}: ')' directly before '}}' (missing ';'?)"
    assert not stack, f"{path.name}: unclosed {stack[-1][0]!r} opened at line {stack[-1][1]}"


def _opener_is_initialiser(toks, close_idx):
    depth = 0
    for j in range(close_idx, -1, -1):
        if toks[j][0] == "op" and toks[j][1] == "}":
            depth += 1
        elif toks[j][0] == "op" and toks[j][1] == "{":
            depth -= 1
            if depth == 0:
                before = toks[j - 1] if j else ("op", "", 0)
                return before[1] in ("=", "]", ",", "{") or before[1] == "("   # array initialiser / annotation value
    return False


@pytest.mark.parametrize("path", FILES, ids=lambda p: str(p.relative_to(ROOT)))
def test_package_and_public_type_match_the_path(path):
    toks = lex(path.read_text())
    words = [t[1] for t in toks]
    assert "package" in words, f"{path}: no package declaration"
    i = words.index("package")
    j = words.index(";", i)
    package = "".join(words[i + 1:j])
    rel = path.relative_to(ROOT).parts
    # java/<flavour>[/<sub-flavour>]/<package path>/<Type>.java: the package path starts at the first com/ or org/
    first = next(k for k, part in enumerate(rel) if part in ("com", "org"))
    assert "/".join(rel[first:-1]) == package.replace(".", "/"), f"{path}: package {package} does not match its directory"
    public_types = [words[k + 2] if words[k + 1] in ("abstract", "final") else words[k + 1 + 0]
                    for k in range(len(words) - 2)
                    if words[k] == "public" and (words[k + 1] in ("class", "interface", "enum")
                                                 or (words[k + 1] in ("abstract", "final") and words[k + 2] in ("class", "interface")))]
    # normalise: the list above holds the keyword for the simple form; pick the name after the keyword
    names = []
    for k in range(len(words) - 1):
        if words[k] in ("class", "interface", "enum") and k > 0 and _is_top_level(toks, k):
            mods = set(words[max(0, k - 3):k])
            if "public" in mods:
                names.append(words[k + 1])
    assert names == [path.stem] or (not names and public_types == []), f"{path}: public top-level types {names} != file name"


def _is_top_level(toks, idx):
    depth = 0
    for kind, text, _ in toks[:idx]:
        if kind == "op" and text == "{":
            depth += 1
        elif kind == "op" and text == "}":
            depth -= 1
    return depth == 0


def _project_classes(flavour_dirs):
    have = set()
    for d in flavour_dirs:
        for p in (ROOT / d).rglob("*.java"):
            rel = p.relative_to(ROOT / d)
            have.add(".".join(rel.with_suffix("").parts))
    return have


FLAVOUR_SOURCES = {
    "hadoop-1.x": ["shared", "hadoop-1.x"],
    "hadoop-1.x-old": ["shared", "hadoop-1.x", "hadoop-1.x-old"],
    "hadoop-2.0.x": ["shared", "hadoop-2.x", "hadoop-2.0.x"],
    "hadoop-2.x": ["shared", "hadoop-2.x"],
}


@pytest.mark.parametrize("flavour", sorted(FLAVOUR_SOURCES))
def test_imports_of_project_classes_resolve_inside_the_flavour(flavour):
    dirs = FLAVOUR_SOURCES[flavour]
    have = _project_classes(dirs)
    for d in dirs:
        for p in (ROOT / d).rglob("*.java"):
            for m in re.finditer(r"^\s*import\s+(static\s+)?(com\.mellanox\.[\w.]+)\s*;", p.read_text(), re.M):
                name = m.group(2)
                if m.group(1):                       # static import: drop the member
                    name = name.rsplit(".", 1)[0]
                if name.endswith(".*"):
                    continue
                assert name in have, f"{p}: import {name} has no source in {dirs}"


@pytest.mark.parametrize("path", FILES, ids=lambda p: str(p.relative_to(ROOT)))
def test_no_duplicate_method_declarations(path):
    """Two methods with the same name and the same parameter types in one class body do not compile."""
    toks = lex(path.read_text())
    sigs = {}
    depth, class_depths = 0, []
    for k, (kind, text, line) in enumerate(toks):
        if kind == "op" and text == "{":
            depth += 1
            # a type body opens when the tokens before '{' contain class/interface/enum since the last ';', '}' or '{'
            j = k - 1
            is_type = False
            while j >= 0 and not (toks[j][0] == "op" and toks[j][1] in ";{}"):
                if toks[j][1] in ("class", "interface", "enum") and (j == 0 or toks[j - 1][1] != "."):
                    is_type = True
                j -= 1
            if is_type:
                class_depths.append(depth)
                sigs[(depth, k)] = {}
        elif kind == "op" and text == "}":
            if class_depths and class_depths[-1] == depth:
                class_depths.pop()
            depth -= 1
        elif kind == "id" and class_depths and depth == class_depths[-1] and k + 1 < len(toks) and toks[k + 1][1] == "(" \
                and k > 0 and toks[k - 1][0] == "id" and toks[k - 1][1] not in ("new", "return", "throw", "else"):
            # `<type> name (` at class-body depth: a method declaration; collect the parameter types up to ')'
            j, par, types, cur = k + 2, 1, [], []
            while j < len(toks) and par:
                t = toks[j][1]
                if t == "(":
                    par += 1
                elif t == ")":
                    par -= 1
                    if par == 0:
                        break
                if t == "," and par == 1:
                    types.append(tuple(cur[:-1])); cur = []
                else:
                    cur.append(t)
                j += 1
            if cur:
                types.append(tuple(x for x in cur[:-1] if x != "final"))
            types = [tuple(x for x in t if x != "final") for t in types]
            body = [key for key in sigs if key[0] == depth][-1]
            sig = (text, tuple(types))
            assert sig not in sigs[body], f"{path.name}:{line}: duplicate declaration of {text}{types}"
            sigs[body][sig] = line
