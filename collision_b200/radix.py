"""Stable LSD radix sort (mirror of collision/radix.py upstream).

RadixSorter keeps the reference's constructor, size rules (check_size),
`num_passes`, `histogram_len`, `resize` and
`sort(cq, keys_buf, out_keys_buf, in_values_buf, out_values_buf, wait_for)`.
The sort itself is csrc/sort.cu: a one-sweep LSD sort over 8-bit digits with
decoupled look-back -- one histogram launch plus one launch per digit --
instead of upstream's block_sort / scan / scatter / copy-back chain per 4-bit
digit (collision/radix.py:140-170).  `group_size` and `radix_bits` are validated
and reported exactly as upstream; they do not shape the launches.  Unlike
upstream, the input buffers are left untouched.
"""
from numpy import dtype

from . import _lib
from .misc import Program, dtype_sizeof, nextPowerOf2, roundUp, same_context
from .misc import np_unsigned_dtypes as _unsigned_names
from .scan import PrefixScanner, PrefixScanProgram  # re-exported like upstream (radix.py:5)

np_unsigned_dtypes = set(map(dtype, _unsigned_names))


class RadixProgram(Program):
    """collision/radix.py:9-25: carries the key/value dtypes of the sorter."""

    def __init__(self, ctx, key_dtype=dtype('uint32'), value_dtype=dtype('uint32')):
        self.key_dtype = dtype(key_dtype)
        self.value_dtype = dtype(value_dtype)
        if self.key_dtype not in np_unsigned_dtypes:
            raise ValueError("Invalid key dtype: {}".format(self.key_dtype))
        if self.key_dtype.itemsize not in (4, 8):
            raise ValueError("Unsupported key dtype: {} (uint32 / uint64)".format(self.key_dtype))
        super().__init__(ctx)
        kb, vb = self.key_dtype.itemsize, dtype_sizeof(self.value_dtype)
        L, P = _lib.lib(), _lib.ptr

        # the reference's kernel-level contracts (collision/radix.py:11-15; tests/test_radix.py:34-53
        # upstream launch them directly): local-memory arguments are placeholders here
        def block_sort(cq, gs, ls, keys, _lk_in, _lk_out, values, _lv_in, _lv_out, histogram, _lh, _count,
                       radix_bits, radix_pass):
            _lib.check(L.col_compat_block_sort(P(keys), P(values), P(histogram), gs // ls, 2 * ls, kb, vb,
                                               int(radix_bits), int(radix_pass), _lib.stream_ptr(cq)),
                       "col_compat_block_sort")

        def scatter(cq, gs, ls, keys, out_keys, values, out_values, offset, _lo, histogram, _lh,
                    radix_bits, radix_pass):
            _lib.check(L.col_compat_scatter(P(keys), P(out_keys), P(values), P(out_values), P(offset), P(histogram),
                                            gs // ls, 2 * ls, kb, vb, int(radix_bits), int(radix_pass),
                                            _lib.stream_ptr(cq)), "col_compat_scatter")

        self.kernels = {'block_sort': _lib.kernel_launcher(block_sort), 'scatter': _lib.kernel_launcher(scatter)}


class RadixSorter:
    histogram_dtype = dtype('uint32')

    def __init__(self, ctx, size, group_size, radix_bits=4,
                 key_dtype=dtype('uint32'), value_dtype=dtype('uint32'),
                 program=None, scan_program=None):
        self.check_size(size, group_size, radix_bits, key_dtype)
        self.size = size
        self.group_size = group_size
        self.radix_bits = radix_bits

        if program is None:
            program = RadixProgram(ctx, key_dtype, value_dtype)
        else:
            if not same_context(program.context, ctx):
                raise ValueError("Sorter and program contexts must match")
            if program.key_dtype != key_dtype:
                raise ValueError("Sorter and program key dtypes must match")
            if program.value_dtype != value_dtype:
                raise ValueError("Sorter and program value dtypes must match")
        self.program = program
        if scan_program is None:
            scan_program = PrefixScanProgram(ctx)
        # kept for API parity (upstream scans the per-pass histogram with it)
        self.scanner = PrefixScanner(ctx, self.histogram_len, self.group_size, scan_program)
        self._scratch = None    # device scratch is allocated on first use

    # -- sizes ------------------------------------------------------------
    @property
    def value_size(self):
        """Bytes per payload slot; a 3-vector travels in a 4-lane slot (radix.py:122-126)."""
        return dtype_sizeof(self.program.value_dtype)

    def _ensure_scratch(self, size=None, cq=None):
        need = _lib.lib().col_sort_scratch_bytes(size or self.size, self.program.key_dtype.itemsize, self.value_size)
        if self._scratch is None or self._scratch.numel() < need:
            self._scratch = _lib.alloc_on(cq, self.program.context, need)
        return self._scratch

    @staticmethod
    def check_size(size, group_size, radix_bits, key_dtype):
        # collision/radix.py:61-74
        key_dtype = dtype(key_dtype)
        if group_size != nextPowerOf2(group_size):
            raise ValueError("Group size ({}) must be a power of two".format(group_size))
        if size % (group_size * 2):
            raise ValueError("Size ({}) must be multiple of 2 * group_size ({})".format(size, group_size))
        if (key_dtype.itemsize * 8) % radix_bits:
            raise ValueError("Radix bits ({}) must evenly divide item-size ({})"
                             .format(radix_bits, key_dtype.itemsize * 8))
        if (2 ** radix_bits) > group_size * 2:
            raise ValueError("2 ^ radix_bits ({}) must be less than 2 * group_size ({})"
                             .format(radix_bits, group_size))

    def resize(self, size=None, group_size=None, radix_bits=None):
        size = self.size if size is None else size
        group_size = self.group_size if group_size is None else group_size
        radix_bits = self.radix_bits if radix_bits is None else radix_bits
        old = (self.size, self.group_size, self.radix_bits)
        self.check_size(size, group_size, radix_bits, self.program.key_dtype)
        self.size, self.group_size, self.radix_bits = size, group_size, radix_bits
        try:
            self.scanner.resize(self.histogram_len, self.group_size)
        except Exception:
            self.size, self.group_size, self.radix_bits = old
            raise

    @property
    def num_passes(self):
        return (self.program.key_dtype.itemsize * 8) // self.radix_bits

    @property
    def histogram_len(self):
        length = (2 ** self.radix_bits) * self.size // 2 // self.group_size
        return roundUp(length, 2 * self.group_size)

    # -- enqueue ----------------------------------------------------------
    def sort(self, cq, keys_buf, out_keys_buf, in_values_buf=None, out_values_buf=None,
             wait_for=None, end_bit=0):
        """Stable sort of `size` keys (and payloads) into the out buffers.
        end_bit (extension): sort on key bits [0, end_bit) only."""
        _lib.begin(cq, wait_for)
        have_values = in_values_buf is not None and out_values_buf is not None
        _lib.check(_lib.lib().col_sort_pairs(
            _lib.ptr(keys_buf), _lib.ptr(out_keys_buf),
            _lib.ptr(in_values_buf) if have_values else None,
            _lib.ptr(out_values_buf) if have_values else None,
            self.size, self.program.key_dtype.itemsize, self.value_size if have_values else 0,
            end_bit, _lib.ptr(self._ensure_scratch(cq=cq)), _lib.stream_ptr(cq)), "col_sort_pairs")
        return _lib.done(cq)

    def radix_pass(self, cq, keys_buf, out_keys_buf, in_values_buf, out_values_buf, radix_pass,
                   wait_for=None):
        """One LSD pass on digit `radix_pass` of width `radix_bits`: what upstream's
        block_sort + prefix_sum + scatter produce per loop iteration (radix.py:140-157)."""
        _lib.begin(cq, wait_for)
        have_values = in_values_buf is not None and out_values_buf is not None
        _lib.check(_lib.lib().col_radix_pass(
            _lib.ptr(keys_buf), _lib.ptr(out_keys_buf),
            _lib.ptr(in_values_buf) if have_values else None,
            _lib.ptr(out_values_buf) if have_values else None,
            self.size, self.program.key_dtype.itemsize, self.value_size if have_values else 0,
            self.radix_bits, radix_pass, _lib.ptr(self._ensure_scratch(cq=cq)), _lib.stream_ptr(cq)), "col_radix_pass")
        return _lib.done(cq)
