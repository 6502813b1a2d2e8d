"""SGF files -> training dataset, drop-in for AlphaGo/preprocessing/game_converter.py:12-218 (the step
before the hot path).

Same class, methods, command line and file contract: `states` uint8 (N, F, size, size), `actions` uint8 (N, 2) as (x, y),
group `file_offsets[name.replace('/', ':')] = [start, count]` (game_converter.py:71-88,122-125); passes are skipped, an
illegal move drops the rest of that game, an unparsable or wrong-size file is skipped with a warning.  What changes is
where the planes come from: positions are replayed on the host rules engine and PACKED (board, stone ages, player, ko:
726 bytes each); the planes of thousands of positions at a time are then produced by one launch of the CUDA encoder
instead of one Python feature pass per position.
"""
import os
import warnings

import numpy as np

from . import go
from .preprocessing import Preprocess, pack_state
from .util import ParseException, sgf_iter_states, sgf_move_list


class SizeMismatchError(Exception):
    pass


class game_converter(object):   # the reference's own (lower-case) class name, game_converter.py:16

    def __init__(self, features, batch=4096):
        self.feature_processor = Preprocess(features)
        self.n_features = self.feature_processor.output_dim
        self.batch = int(batch)   # positions per encoder launch in the bulk path

    # ------------------------------------------------------------------ reference-shaped generator
    def convert_game(self, file_name, bd_size):
        """Read the given SGF file into an iterable of (input, output) pairs: input = (1, F, size, size) one-hot planes of
        the position, output = the move played there as (x, y); passes are skipped; SizeMismatchError if the game is not
        played on bd_size (game_converter.py:22-40)"""
        with open(file_name, "r") as file_object:
            state_action_iterator = sgf_iter_states(file_object.read())
        for (state, move, player) in state_action_iterator:
            if state.size != bd_size:
                raise SizeMismatchError()
            if move != go.PASS_MOVE:
                yield (self.feature_processor.state_to_tensor(state), move)

    # ------------------------------------------------------------------ bulk path
    def pack_game(self, file_name, bd_size):
        """The same positions as convert_game, packed for the encoder instead of encoded: returns (packed, actions, error)
        with packed = list of (board int8[n], age uint8[n], player, ko), actions = list of (x, y), error = the exception
        that ended the game early (None if it ran to its end).  Pairs produced before an error are kept, as the
        reference keeps what it had written (game_converter.py:103-123)."""
        packed, actions, error = [], [], None
        try:
            with open(file_name, "r") as file_object:
                text = file_object.read()
            gs, moves, players, parse_error = sgf_move_list(text)
            if moves and gs.size != bd_size:
                raise SizeMismatchError()
            # one library call replays the record and packs the position in front of every non-pass move
            b, a, pl, ko, illegal = gs.replay_pack(moves, players)
            packed = [(b[i], a[i], int(pl[i]), int(ko[i])) for i in range(len(b))]
            played = moves if illegal is None else moves[:illegal + 1]
            actions = [m for m in played if m is not go.PASS_MOVE]
            if illegal is not None:
                error = go.IllegalMove(str(moves[illegal]))
            elif parse_error is not None:   # a bad node part-way through: the pairs in front of it stand
                error = parse_error
        except Exception as e:   # classified by the caller
            error = e
        return packed, actions, error

    def encode_packed_positions(self, packed, bd_size, bits=False):
        """list of packed positions -> uint8 planes (N, F, size, size) on the host, `self.batch` positions per launch.
        bits=True: each position's planes come back bit-packed (np.packbits order, (N, ceil(F*size*size/8))): an eighth
        of the device->host traffic and of what the archive writer has to compress."""
        import torch
        n = len(packed)
        row = self.n_features * bd_size * bd_size
        out = np.empty((n, (row + 7) // 8), np.uint8) if bits else np.empty((n, self.n_features, bd_size, bd_size), np.uint8)
        for s in range(0, n, self.batch):
            chunk = packed[s:s + self.batch]
            m = len(chunk)
            board = torch.from_numpy(np.stack([p[0] for p in chunk])).cuda(non_blocking=True)
            age = torch.from_numpy(np.stack([p[1] for p in chunk])).cuda(non_blocking=True)
            player = torch.from_numpy(np.asarray([p[2] for p in chunk], np.int8)).cuda(non_blocking=True)
            ko = torch.from_numpy(np.asarray([p[3] for p in chunk], np.int16)).cuda(non_blocking=True)
            planes = self.feature_processor.encode_packed(board, age, player, ko, bd_size)
            if bits:
                flat = planes.reshape(m, row)
                if row % 8:
                    flat = torch.nn.functional.pad(flat, (0, 8 - row % 8))
                w = torch.tensor([128, 64, 32, 16, 8, 4, 2, 1], dtype=torch.int32, device=flat.device)
                out[s:s + m] = (flat.reshape(m, -1, 8).to(torch.int32) * w).sum(-1).to(torch.uint8).cpu().numpy()
            else:
                out[s:s + m] = planes.cpu().numpy()
        return out

    def sgfs_to_hdf5(self, sgf_files, hdf5_file, bd_size=19, ignore_errors=True, verbose=False, container="hdf5"):
        """Convert all files in the iterable sgf_files into one dataset file (game_converter.py:42-140).

        The resulting file has `states` (n_data, n_features, size, size), `actions` (n_data, 2) and the group
        `file_offsets` mapping file names ('/' -> ':') to [index, length]; it is written under a hidden temporary name
        and renamed on success.  container="hdf5" (default) is the reference's HDF5 layout -- extendable datasets in
        compressed chunks of 64 / 1024 rows -- written through h5py where it is importable, else through
        hdf5_lite's writer; each encoded block is appended as it comes off the device, so memory holds one block, not
        the dataset.  container="npz" writes the same three members as an .npz archive with the planes bit-packed
        (smaller and faster to write; supervised_policy_trainer.open_dataset reads either)."""
        if container not in ("hdf5", "npz"):
            raise ValueError("container must be 'hdf5' or 'npz'")
        tmp_file = os.path.join(os.path.dirname(hdf5_file), ".tmp." + os.path.basename(hdf5_file))
        names, offsets = [], []
        pending, all_actions, blocks = [], [], []
        next_idx = 0
        sink = _Hdf5Sink(tmp_file, self.n_features, bd_size) if container == "hdf5" else None
        try:
            for file_name in sgf_files:
                if verbose:
                    print(file_name)
                packed, actions, error = self.pack_game(file_name, bd_size)
                if error is not None:
                    if isinstance(error, go.IllegalMove):
                        warnings.warn("Illegal Move encountered in %s\n\tdropping the remainder of the game" % file_name)
                    elif isinstance(error, ParseException):
                        warnings.warn("Could not parse %s\n\tdropping game" % file_name)
                    elif isinstance(error, SizeMismatchError):
                        warnings.warn("Skipping %s; wrong board size" % file_name)
                    elif ignore_errors:
                        warnings.warn("Unkown exception with file %s\n\t%s" % (file_name, error))
                    else:
                        raise error
                n_pairs = len(packed)
                if n_pairs > 0:
                    names.append(file_name.replace("/", ":"))   # '/' has special meaning in HDF5 key names
                    offsets.append([next_idx, n_pairs])
                    pending.extend(packed)
                    all_actions.extend(actions)
                    next_idx += n_pairs
                    if verbose:
                        print("\t%d state/action pairs extracted" % n_pairs)
                elif verbose:
                    print("\t-no usable data-")
                if len(pending) >= self.batch:
                    block = self.encode_packed_positions(pending, bd_size, bits=True)   # bit-packed over the bus either way
                    pending = []
                    if sink is not None:
                        sink.append_states(block)
                    else:
                        blocks.append(block)
            if pending:
                block = self.encode_packed_positions(pending, bd_size, bits=True)
                if sink is not None:
                    sink.append_states(block)
                else:
                    blocks.append(block)
            shape = (next_idx, self.n_features, bd_size, bd_size)
            acts = np.asarray(all_actions, np.uint8).reshape(-1, 2)
            offs = np.asarray(offsets, np.int64).reshape(-1, 2)
            if sink is not None:
                sink.finish(acts, names, offs)
            else:
                states = np.concatenate(blocks) if blocks else np.zeros((0, (shape[1] * bd_size * bd_size + 7) // 8), np.uint8)
                _write_archive(tmp_file, states, shape, acts, names, offs)
        except Exception:
            print("sgfs_to_hdf5 failed")
            if sink is not None:
                sink.abandon()
            if os.path.exists(tmp_file):
                os.remove(tmp_file)
            raise
        if verbose:
            print("finished. renaming %s to %s" % (tmp_file, hdf5_file))
        os.rename(tmp_file, hdf5_file)


class _Hdf5Sink(object):
    """the dataset file while it grows: the reference's three members (game_converter.py:71-88), filled block by block
    through h5py's resize-and-assign calls -- h5py itself where importable, hdf5_lite's writer (same calls) otherwise.
    Blocks are unpacked, compressed and written by one background thread (the codecs and numpy release the interpreter lock),
    so the file keeps up while the caller replays the next games."""

    def __init__(self, path, n_features, size):
        from concurrent.futures import ThreadPoolExecutor
        from . import hdf5_lite
        self.row = (n_features, size, size)
        self.f = hdf5_lite.open_file(path, "w")
        self.states = self.f.create_dataset("states", shape=(0,) + self.row, dtype=np.uint8, maxshape=(None,) + self.row,
                                            chunks=(64,) + self.row, compression="lzf")
        self.n = 0
        self.worker = ThreadPoolExecutor(max_workers=1)   # one thread: blocks stay in order
        self.jobs = []

    def _write(self, bits):
        m = len(bits)
        planes = np.unpackbits(bits, axis=1)[:, :int(np.prod(self.row))].reshape((m,) + self.row)
        self.states.resize(self.n + m, axis=0)
        self.states[self.n:self.n + m] = planes
        self.n += m

    def append_states(self, bits):
        """rows of bit-packed planes (np.packbits order) as they come back from the device"""
        while len(self.jobs) > 2:            # bounded backlog: at most three blocks in memory
            self.jobs.pop(0).result()
        self.jobs.append(self.worker.submit(self._write, bits))

    def finish(self, actions, names, offsets):
        for j in self.jobs:
            j.result()                        # re-raises what the writer thread hit
        self.worker.shutdown()
        assert len(actions) == self.n
        self.f.create_dataset("actions", data=actions, dtype=np.uint8, maxshape=(None, 2), chunks=(1024, 2), compression="lzf")
        g = self.f.require_group("file_offsets")
        for name, off in zip(names, offsets):
            g[name] = [int(off[0]), int(off[1])]
        self.f.close()

    def abandon(self):
        self.worker.shutdown(wait=True, cancel_futures=True)
        try:
            self.f.close()
        except Exception:
            pass


def _write_archive(path, states, shape, actions, names, offsets):
    """an .npz archive (np.load reads it), members written one by one with the fastest deflate level; the planes are stored
    bit-packed (`states_bits` + `states_shape`; supervised_policy_trainer.open_dataset unpacks them)"""
    import zipfile
    members = {"states_bits": states, "states_shape": np.asarray(shape, np.int64), "actions": actions,
               "file_offsets_names": np.asarray(names), "file_offsets": offsets}
    with zipfile.ZipFile(path, "w", compression=zipfile.ZIP_DEFLATED, compresslevel=1, allowZip64=True) as z:
        for name, arr in members.items():
            with z.open(name + ".npy", "w", force_zip64=True) as f:
                np.lib.format.write_array(f, np.ascontiguousarray(arr), allow_pickle=False)


def run_game_converter(cmd_line_args=None):
    """Run conversions. command-line args may be passed in as a list (game_converter.py:143-214)"""
    import argparse
    import sys

    parser = argparse.ArgumentParser(
        description="Prepare SGF Go game files for training the neural network model.",
        epilog="Available features are: board, ones, turns_since, liberties, capture_size, self_atari_size, "
               "liberties_after, ladder_capture, ladder_escape, sensibleness, and zeros.")
    parser.add_argument("--features", "-f", help="Comma-separated list of features to compute and store or 'all'", default="all")
    parser.add_argument("--outfile", "-o", help="Destination to write data (hdf5 file)", required=True)
    parser.add_argument("--recurse", "-R", help="Set to recurse through directories searching for SGF files", default=False, action="store_true")
    parser.add_argument("--directory", "-d", help="Directory containing SGF files to process. if not present, expects files from stdin", default=None)
    parser.add_argument("--size", "-s", help="Size of the game board. SGFs not matching this are discarded with a warning", type=int, default=19)
    parser.add_argument("--verbose", "-v", help="Turn on verbose mode", default=False, action="store_true")
    parser.add_argument("--container", help="File format of --outfile: hdf5 (the reference's layout; default) or npz (bit-packed archive, "
                        "faster to write; the trainer reads both)", choices=["hdf5", "npz"], default="hdf5")   # not in the reference
    args = parser.parse_args(cmd_line_args)

    if args.features.lower() == "all":
        # the reference's 'all' leaves the two ladder features out because it cannot compute them
        # (game_converter.py:166-177); the list is kept so that files stay interchangeable -- ask for them by name
        feature_list = ["board", "ones", "turns_since", "liberties", "capture_size", "self_atari_size", "liberties_after",
                        "sensibleness", "zeros"]
    else:
        feature_list = args.features.split(",")
    if args.verbose:
        print("using features", feature_list)

    converter = game_converter(feature_list)

    def _is_sgf(fname):
        return fname.strip()[-4:] == ".sgf"

    def _walk_all_sgfs(root):
        for (dirpath, dirname, files) in os.walk(root):
            for filename in files:
                if _is_sgf(filename):
                    yield os.path.join(dirpath, filename)

    def _list_sgfs(path):
        return (os.path.join(path, f) for f in os.listdir(path) if _is_sgf(f))

    if args.directory:
        files = _walk_all_sgfs(args.directory) if args.recurse else _list_sgfs(args.directory)
    else:
        files = (f.strip() for f in sys.stdin if _is_sgf(f))
    converter.sgfs_to_hdf5(files, args.outfile, bd_size=args.size, verbose=args.verbose, container=args.container)


if __name__ == "__main__":
    run_game_converter()
