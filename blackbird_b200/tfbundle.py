"""TensorFlow checkpoint bundles without TensorFlow: read and write the `best.index` /
`best.data-00000-of-00001` pair that tf.compat.v1.train.Saver().save(sess, 'blackbird_models/<Name>_<Version>/best')
produces (Network.py:86-99), so that weights trained by the reference can drive this engine and weights trained here
can be restored by the reference's own graph.

Format (TensorFlow's tensor bundle, "checkpoint V2"; restated from the published format, no TensorFlow code involved):

* `<prefix>.data-00000-of-00001` -- the raw little-endian bytes of every tensor, back to back.
* `<prefix>.index` -- an immutable sorted string table in LevelDB's table format: data blocks of prefix-compressed
  (shared, non_shared, value_len varints + key suffix + value) entries with a restart array, each block followed by a
  1-byte compression type (0) and a masked CRC32C; a (here empty) metaindex block; an index block mapping the last key
  of every data block to its (offset, size) handle; a 48-byte footer ending in the magic 0xdb4775248b80fb57.  Key ""
  holds a BundleHeaderProto {num_shards = 1, version.producer = 1}; every other key is a variable name whose value is
  a BundleEntryProto {dtype = 1 (DT_FLOAT), shape, shard_id = 0, offset, size, crc32c (masked, of the tensor bytes)}.
* `<prefix>.meta` (the MetaGraphDef that Network.loadModel feeds to import_meta_graph, Network.py:101-112) is NOT
  produced: it is a serialised TensorFlow graph.  A bundle written here is restored by the reference into the graph
  its own NetworkFactory builds (or next to a `best.meta` the reference saved for the same configuration).

Variable names are the reference graph's (NetworkFactory.py:40-172 variable scopes, tf.compat.v1.layers naming):
resTower/conv_block/{conv,batch_norm}, resTower/block_<i>/{conv_1,batch_norm_1,conv_2,batch_norm_2},
value/{convolution,batch_norm,dense_1,dense_2}, policy/{convolution,batch_norm,policy}; kernel / bias,
gamma / beta / moving_mean / moving_variance.  Optimizer slot variables found in a bundle are ignored on reading.

PARITY UNPINNED: TensorFlow cannot be installed here, so no bundle written by the reference is available to read
back; the format restatement is checked against its own reader, LevelDB's CRC32C known answers and the footer magic.
"""
import os
import struct

import numpy as np

MAGIC = 0xdb4775248b80fb57
_MASK_DELTA = 0xa282ead8


# ---- CRC32C (Castagnoli), table driven -------------------------------------------------------------------------------
def _make_table():
    t = np.zeros(256, np.uint32)
    for i in range(256):
        c = i
        for _ in range(8):
            c = (c >> 1) ^ 0x82F63B78 if c & 1 else c >> 1
        t[i] = c
    return t


_TABLE = _make_table()


def crc32c(data):
    crc = 0xFFFFFFFF
    tbl = _TABLE
    for b in bytes(data):
        crc = int(tbl[(crc ^ b) & 0xFF]) ^ (crc >> 8)
    return crc ^ 0xFFFFFFFF


def crc32c_np(arr):
    """CRC32C of a (possibly large) buffer, 8 bytes at a time in numpy where possible."""
    data = np.frombuffer(bytes(arr), np.uint8)
    crc = 0xFFFFFFFF
    # plain table loop in chunks: tensors here are at most a few hundred KB
    for b in data.tolist():
        crc = int(_TABLE[(crc ^ b) & 0xFF]) ^ (crc >> 8)
    return crc ^ 0xFFFFFFFF


def mask_crc(crc):
    return (((crc >> 15) | (crc << 17)) + _MASK_DELTA) & 0xFFFFFFFF


def unmask_crc(masked):
    rot = (masked - _MASK_DELTA) & 0xFFFFFFFF
    return ((rot >> 17) | (rot << 15)) & 0xFFFFFFFF


# ---- varints / tiny protobuf -----------------------------------------------------------------------------------------
def _varint(n):
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        out.append(b | 0x80 if n else b)
        if not n:
            return bytes(out)


def _read_varint(buf, i):
    n, shift = 0, 0
    while True:
        b = buf[i]
        i += 1
        n |= (b & 0x7F) << shift
        shift += 7
        if not b & 0x80:
            return n, i


def _parse_proto(buf):
    """-> list of (field number, wire type, value) for varint (0), fixed64 (1), bytes (2) and fixed32 (5) fields"""
    out, i, buf = [], 0, bytes(buf)
    while i < len(buf):
        tag, i = _read_varint(buf, i)
        num, wire = tag >> 3, tag & 7
        if wire == 0:
            v, i = _read_varint(buf, i)
        elif wire == 1:
            v = struct.unpack_from('<Q', buf, i)[0]
            i += 8
        elif wire == 2:
            n, i = _read_varint(buf, i)
            v = buf[i:i + n]
            i += n
        elif wire == 5:
            v = struct.unpack_from('<I', buf, i)[0]
            i += 4
        else:
            raise ValueError('unsupported protobuf wire type %d' % wire)
        out.append((num, wire, v))
    return out


def _entry_proto(shape, offset, size, crc_masked):
    """BundleEntryProto{dtype=1 DT_FLOAT, shape=2, shard_id=3, offset=4, size=5, crc32c=6 (fixed32)}; proto3 zero fields omitted."""
    dims = b''.join(b'\x12' + _varint(len(d)) + d for d in (b'\x08' + _varint(int(s)) for s in shape))
    out = b'\x08\x01' + b'\x12' + _varint(len(dims)) + dims
    if offset:
        out += b'\x20' + _varint(offset)
    out += b'\x28' + _varint(size) + b'\x35' + struct.pack('<I', crc_masked)
    return out


_HEADER_PROTO = b'\x08\x01\x1a\x02\x08\x01'      # BundleHeaderProto{num_shards: 1, version{producer: 1}} (endianness LITTLE = 0 omitted)


# ---- table writer / reader ---------------------------------------------------------------------------------------------
def _build_block(items, restart_interval=16):
    out, restarts, last = bytearray(), [], b''
    for n, (k, v) in enumerate(items):
        shared = 0
        if n % restart_interval == 0:
            restarts.append(len(out))
        else:
            m = min(len(last), len(k))
            while shared < m and last[shared] == k[shared]:
                shared += 1
        out += _varint(shared) + _varint(len(k) - shared) + _varint(len(v)) + k[shared:] + v
        last = k
    if not restarts:
        restarts = [0]
    for r in restarts:
        out += struct.pack('<I', r)
    out += struct.pack('<I', len(restarts))
    return bytes(out)


def _with_trailer(block):
    return block + b'\x00' + struct.pack('<I', mask_crc(crc32c(block + b'\x00')))


def _parse_block(buf):
    buf = bytes(buf)
    n_restarts = struct.unpack_from('<I', buf, len(buf) - 4)[0]
    end = len(buf) - 4 - 4 * n_restarts
    items, i, last = [], 0, b''
    while i < end:
        shared, i = _read_varint(buf, i)
        non_shared, i = _read_varint(buf, i)
        vlen, i = _read_varint(buf, i)
        k = last[:shared] + buf[i:i + non_shared]
        i += non_shared
        items.append((k, buf[i:i + vlen]))
        i += vlen
        last = k
    return items


def write_bundle(prefix, tensors):
    """tensors: dict name -> float32 ndarray.  Writes <prefix>.index and <prefix>.data-00000-of-00001."""
    os.makedirs(os.path.dirname(prefix) or '.', exist_ok=True)
    names = sorted(tensors, key=lambda s: s.encode())
    data, items = bytearray(), [(b'', _HEADER_PROTO)]
    for name in names:
        arr = np.ascontiguousarray(tensors[name], dtype='<f4')
        raw = arr.tobytes()
        items.append((name.encode(), _entry_proto(arr.shape, len(data), len(raw), mask_crc(crc32c(raw)))))
        data += raw
    with open(prefix + '.data-00000-of-00001', 'wb') as f:
        f.write(bytes(data))
    block = _with_trailer(_build_block(items))
    meta = _with_trailer(_build_block([]))
    index_items = [(items[-1][0], _varint(0) + _varint(len(block) - 5))]
    index = _with_trailer(_build_block(index_items, restart_interval=1))
    off_meta, off_index = len(block), len(block) + len(meta)
    footer = _varint(off_meta) + _varint(len(meta) - 5) + _varint(off_index) + _varint(len(index) - 5)
    footer += b'\x00' * (40 - len(footer)) + struct.pack('<Q', MAGIC)
    with open(prefix + '.index', 'wb') as f:
        f.write(block + meta + index + footer)


def read_bundle(prefix, check_crc=True):
    """-> dict name -> ndarray for every DT_FLOAT (1), DT_DOUBLE (2), DT_INT32 (3) or DT_INT64 (9) tensor of the bundle."""
    idx = open(prefix + '.index', 'rb').read()
    if len(idx) < 48 or struct.unpack_from('<Q', idx, len(idx) - 8)[0] != MAGIC:
        raise ValueError('%s.index is not a TensorFlow bundle index (bad footer magic)' % prefix)
    footer = idx[-48:]
    _, i = _read_varint(footer, 0)
    _, i = _read_varint(footer, i)
    ioff, i = _read_varint(footer, i)
    isize, i = _read_varint(footer, i)

    def block(off, size):
        body = idx[off:off + size]
        if idx[off + size] != 0:
            raise ValueError('compressed index blocks are not supported')
        if check_crc and unmask_crc(struct.unpack_from('<I', idx, off + size + 1)[0]) != crc32c(idx[off:off + size + 1]):
            raise ValueError('index block checksum mismatch')
        return _parse_block(body)

    entries = []
    for _, handle in block(ioff, isize):
        off, j = _read_varint(handle, 0)
        size, j = _read_varint(handle, j)
        entries += block(off, size)
    shards = {}
    out = {}
    dtypes = {1: '<f4', 2: '<f8', 3: '<i4', 9: '<i8'}
    for key, val in entries:
        if key == b'':
            continue
        f = {num: v for num, _, v in _parse_proto(val)}
        dims = [dict((n, v) for n, _, v in _parse_proto(d)).get(1, 0) for n_, w_, d in _parse_proto(f.get(2, b'')) if n_ == 2]
        dt = dtypes.get(f.get(1, 0))
        if dt is None:
            continue
        shard = f.get(3, 0)
        if shard not in shards:
            n_shards = 1
            for k0, v0 in entries:
                if k0 == b'':
                    n_shards = {num: v for num, _, v in _parse_proto(v0)}.get(1, 1)
            shards[shard] = open('%s.data-%05d-of-%05d' % (prefix, shard, n_shards), 'rb').read()
        off, size = f.get(4, 0), f.get(5, 0)
        raw = shards[shard][off:off + size]
        if check_crc and 6 in f and unmask_crc(f[6]) != crc32c(raw):
            raise ValueError('tensor %r: checksum mismatch' % key.decode())
        out[key.decode()] = np.frombuffer(raw, dtype=dt).reshape(dims).copy()
    return out


# ---- the reference graph's variables <-> the engine's flat weight vector -------------------------------------------------
def variable_layout(cin, actions, blocks, filters, dense):
    """[(TF variable name, shape)] in the engine's flat order (DESIGN.md "weight layout")."""
    out = []

    def conv_bn(conv, bn, k, ci, co):
        out.append((conv + '/kernel', (k, k, ci, co)))
        out.append((conv + '/bias', (co,)))
        for v in ('gamma', 'beta', 'moving_mean', 'moving_variance'):
            out.append((bn + '/' + v, (co,)))

    conv_bn('resTower/conv_block/conv', 'resTower/conv_block/batch_norm', 3, cin, filters)
    for b in range(blocks):
        conv_bn('resTower/block_%d/conv_1' % b, 'resTower/block_%d/batch_norm_1' % b, 3, filters, filters)
        conv_bn('resTower/block_%d/conv_2' % b, 'resTower/block_%d/batch_norm_2' % b, 3, filters, filters)
    conv_bn('value/convolution', 'value/batch_norm', 1, filters, 1)
    out += [('value/dense_1/kernel', (1, dense)), ('value/dense_1/bias', (dense,)),
            ('value/dense_2/kernel', (dense, 1)), ('value/dense_2/bias', (1,))]
    conv_bn('policy/convolution', 'policy/batch_norm', 1, filters, 2)
    out += [('policy/policy/kernel', (2, actions)), ('policy/policy/bias', (actions,))]
    return out


def flat_to_tensors(flat, cin, actions, blocks, filters, dense):
    flat = np.asarray(flat, np.float32)
    out, o = {}, 0
    for name, shape in variable_layout(cin, actions, blocks, filters, dense):
        n = int(np.prod(shape))
        out[name] = flat[o:o + n].reshape(shape)
        o += n
    if o != flat.size:
        raise ValueError('flat weight vector has %d floats, the shape needs %d' % (flat.size, o))
    return out


def tensors_to_flat(tensors):
    """Flat vector + (cin, actions, blocks, filters, dense) recovered from the variables' shapes."""
    k0 = tensors['resTower/conv_block/conv/kernel']
    cin, filters = int(k0.shape[2]), int(k0.shape[3])
    blocks = 0
    while 'resTower/block_%d/conv_1/kernel' % blocks in tensors:
        blocks += 1
    dense = int(tensors['value/dense_1/kernel'].shape[1])
    actions = int(tensors['policy/policy/kernel'].shape[1])
    parts = []
    for name, shape in variable_layout(cin, actions, blocks, filters, dense):
        t = np.asarray(tensors[name], np.float32)
        if tuple(t.shape) != tuple(shape):
            raise ValueError('variable %s has shape %s, expected %s' % (name, t.shape, shape))
        parts.append(t.ravel())
    return np.concatenate(parts), (cin, actions, blocks, filters, dense)
