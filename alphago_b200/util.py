"""SGF helpers with the surface of the reference `AlphaGo/util.py` (util.py:6-83): `flatten_idx`, `unflatten_idx`,
`sgf_to_gamestate`, `sgf_iter_states`.  The reference parses with the third-party `sgf==0.5` package
(requirements.txt), which this image does not have; what it needs from it is the main line of the first game --
`collection[0].root.properties` and `.rest` -- so a small reader of exactly that is included here (FF[4] grammar:
game trees, nodes, properties with `\\` escapes inside values; at every branching point the first variation is the
main line).
"""
import re
import string

from . import go

_LETTERS = string.ascii_lowercase + string.ascii_uppercase   # Python 2's string.letters in the C locale (util.py:20-21)


class ParseException(Exception):
    """what `sgf.ParseException` is to the reference (game_converter.py:114): the text is not an SGF collection"""


def flatten_idx(position, size):
    (x, y) = position
    return x * size + y


def unflatten_idx(idx, size):
    x, y = divmod(idx, size)
    return (x, y)


# ------------------------------------------------------------------------------------------ SGF main line
_WS = re.compile(r"\s*")
_IDENT = re.compile(r"[^\W\d_]+")                 # letters, as str.isalpha
_PLAIN_VALUE = re.compile(r"\[([^\]\\]*)\]")     # a value without escapes: the common case, matched in one step


def _skip_ws(s, i):
    return _WS.match(s, i).end()


def _parse_value(s, i):
    """s[i] == '[': returns (text, index after the closing bracket); a backslash keeps the next character"""
    out = []
    i += 1
    n = len(s)
    while True:
        if i >= n:
            raise ParseException("unterminated property value")
        c = s[i]
        if c == "\\":
            if i + 1 >= n:
                raise ParseException("dangling escape")
            nxt = s[i + 1]
            if nxt == "\n" or nxt == "\r":       # soft line break: removed
                i += 2
                if i < n and s[i] in "\r\n" and s[i] != nxt:
                    i += 1
                continue
            out.append(nxt)
            i += 2
            continue
        if c == "]":
            return "".join(out), i + 1
        out.append(c)
        i += 1


def _parse_node(s, i):
    """s[i] == ';': returns ({ident: [values]}, index after the node)"""
    props = {}
    i = _skip_ws(s, i + 1)
    n = len(s)
    while True:
        m = _IDENT.match(s, i)
        if m is None:
            break
        ident = m.group()
        if not ident.isupper():
            ident = "".join(ch for ch in ident if ch.isupper())   # old files write e.g. "SiZe": only capitals count
        i = _skip_ws(s, m.end())
        if i >= n or s[i] != "[":
            raise ParseException("property %r has no value" % ident)
        values = props.setdefault(ident, [])
        while i < n and s[i] == "[":
            m = _PLAIN_VALUE.match(s, i)
            if m is not None:
                v, i = m.group(1), m.end()
            else:
                v, i = _parse_value(s, i)   # escapes, or an unterminated value (raises)
            values.append(v)
            i = _skip_ws(s, i)
    return props, i


def sgf_main_line(sgf_string):
    """nodes (property dicts) of the main line of the first game in the collection: root first"""
    s = sgf_string
    i = s.find("(")
    if i < 0:
        raise ParseException("no game tree")
    nodes = []
    depth_main = True
    i = _skip_ws(s, i + 1)
    n = len(s)
    while i < n:
        c = s[i]
        if c == ";":
            props, i = _parse_node(s, i)
            nodes.append(props)
        elif c == "(":        # first variation at this point continues the main line
            i += 1
        elif c == ")":        # the main line's leaf: everything after it is a side variation or another game
            break
        elif c.isspace():
            i += 1
        else:
            raise ParseException("unexpected %r at offset %d" % (c, i))
    if not nodes or not depth_main:
        raise ParseException("empty game tree")
    return nodes


def _parse_sgf_move(node_value):
    """Given a well-formed move string, return either PASS_MOVE or the (x, y) position (util.py:16-26)"""
    if node_value == "" or node_value == "tt":
        return go.PASS_MOVE
    row = _LETTERS.index(node_value[1])
    col = _LETTERS.index(node_value[0])
    return (col, row)   # GameState expects (x, y) where x is column and y is row


def _sgf_init_gamestate(root_props):
    """GameState from the root node (util.py:29-49): SZ, AB/AW stones played through do_move (so they enter the history
    and can capture), then PL"""
    s_size = root_props.get("SZ", ["19"])[0]
    s_player = root_props.get("PL", ["B"])[0]
    gs = go.GameState(int(s_size))
    for stone in root_props.get("AB", []):
        gs.do_move(_parse_sgf_move(stone), go.BLACK)
    for stone in root_props.get("AW", []):
        gs.do_move(_parse_sgf_move(stone), go.WHITE)
    gs.current_player = go.BLACK if s_player == "B" else go.WHITE
    return gs


def sgf_iter_states(sgf_string):
    """Iterates over (GameState, move, player) tuples in the first game of the given SGF text, main line only.  As in
    the reference (util.py:64-83) the SAME GameState object is yielded every time and advanced in place after each
    yield, a node carrying both colours counts as White's, and a node without a move repeats the previous one."""
    nodes = sgf_main_line(sgf_string)
    gs = _sgf_init_gamestate(nodes[0])
    have = False
    move = player = None
    for props in nodes[1:]:
        if "W" in props:
            move, player, have = _parse_sgf_move(props["W"][0]), go.WHITE, True
        elif "B" in props:
            move, player, have = _parse_sgf_move(props["B"][0]), go.BLACK, True
        if not have:
            raise UnboundLocalError("local variable 'move' referenced before assignment")   # what util.py:81 does
        yield (gs, move, player)
        gs.do_move(move, player)


def sgf_move_list(sgf_string):
    """(GameState after the root node's set-up, moves, players, error): what sgf_iter_states would play, as lists, for
    callers that replay a whole record in one library call (GameState.replay_pack).  Same node rules as sgf_iter_states.
    `error` is the exception a node raised part-way through the record (a malformed move value, a first node with no
    move), or None: the moves in front of it are returned, because the reference's generator has already yielded -- and
    its converter already written -- those pairs when the exception surfaces (game_converter.py:103-123)."""
    nodes = sgf_main_line(sgf_string)
    gs = _sgf_init_gamestate(nodes[0])
    moves, players = [], []
    have = False
    move = player = None
    try:
        for props in nodes[1:]:
            if "W" in props:
                move, player, have = _parse_sgf_move(props["W"][0]), go.WHITE, True
            elif "B" in props:
                move, player, have = _parse_sgf_move(props["B"][0]), go.BLACK, True
            if not have:
                raise UnboundLocalError("local variable 'move' referenced before assignment")   # what util.py:81 does
            moves.append(move)
            players.append(player)
    except Exception as e:
        return gs, moves, players, e
    return gs, moves, players, None


def sgf_to_gamestate(sgf_string):
    """GameState at the end of the first game's main line (util.py:52-61)"""
    gs = None
    for (gs, move, player) in sgf_iter_states(sgf_string):
        pass
    if gs is None:   # no move nodes: the set-up position
        gs = _sgf_init_gamestate(sgf_main_line(sgf_string)[0])
    return gs
