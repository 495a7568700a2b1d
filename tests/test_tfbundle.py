"""CPU: the TensorFlow checkpoint-bundle reader / writer (blackbird_b200/tfbundle.py): CRC32C known
answers, LevelDB table structure, round trips through the reference graph's variable names."""
import os
import struct

import numpy as np
import pytest

from blackbird_b200 import netinit, tfbundle


def test_crc32c_known_answers():
    assert tfbundle.crc32c(b'123456789') == 0xE3069283                 # the standard CRC-32C check value
    assert tfbundle.crc32c(b'\x00' * 32) == 0x8A9136AA                 # RFC 3720 B.4
    assert tfbundle.crc32c(b'\xff' * 32) == 0x62A8AB43
    assert tfbundle.crc32c(bytes(range(32))) == 0x46DD794E
    for v in (0, 1, 0xE3069283, 0xFFFFFFFF):
        assert tfbundle.unmask_crc(tfbundle.mask_crc(v)) == v
    assert tfbundle.mask_crc(tfbundle.crc32c(b'foo')) != tfbundle.crc32c(b'foo')


@pytest.mark.parametrize('shape', [(3, 7, 4, 16, 16), (17, 4032, 2, 16, 8), (3, 9, 0, 8, 5)])
def test_round_trip_through_reference_variable_names(tmp_path, shape):
    cin, A, blocks, filters, dense = shape
    flat = netinit.random_weights(cin, A, blocks=blocks, filters=filters, dense=dense, seed=3)
    flat = flat + np.random.RandomState(0).standard_normal(flat.size).astype(np.float32) * 0.01
    tensors = tfbundle.flat_to_tensors(flat, cin, A, blocks, filters, dense)
    assert tensors['resTower/conv_block/conv/kernel'].shape == (3, 3, cin, filters)
    assert tensors['policy/policy/kernel'].shape == (2, A) and tensors['value/dense_1/kernel'].shape == (1, dense)
    assert len(tensors) == 6 * (2 * blocks + 3) + 6
    prefix = str(tmp_path / 'blackbird_models' / 'm_1' / 'best')
    tfbundle.write_bundle(prefix, tensors)
    back = tfbundle.read_bundle(prefix)
    assert sorted(back) == sorted(tensors)
    flat2, got_shape = tfbundle.tensors_to_flat(back)
    assert got_shape == shape and np.array_equal(flat2, flat)
    # structure: footer magic, one data file of exactly the tensors' bytes, sorted keys with the header first
    idx = open(prefix + '.index', 'rb').read()
    assert struct.unpack_from('<Q', idx, len(idx) - 8)[0] == 0xdb4775248b80fb57
    assert os.path.getsize(prefix + '.data-00000-of-00001') == flat.size * 4
    # corruption is detected
    data = bytearray(open(prefix + '.data-00000-of-00001', 'rb').read())
    data[100] ^= 1
    open(prefix + '.data-00000-of-00001', 'wb').write(bytes(data))
    with pytest.raises(ValueError):
        tfbundle.read_bundle(prefix)
    assert tfbundle.read_bundle(prefix, check_crc=False)


def test_reader_handles_several_data_blocks_and_extra_variables(tmp_path):
    """A Saver bundle of the reference also holds optimizer slots and may spread its entries over several index blocks."""
    rs = np.random.RandomState(1)
    names = ['training/beta1_power'] + ['resTower/block_%d/conv_1/kernel/Adam' % i for i in range(40)] + ['value/dense_2/bias']
    tensors = {n: rs.standard_normal((3, 2)).astype(np.float32) for n in names}
    prefix = str(tmp_path / 'best')
    # hand-built index with three data blocks
    items, data = [(b'', tfbundle._HEADER_PROTO)], bytearray()
    for n in sorted(tensors, key=lambda s: s.encode()):
        raw = tensors[n].tobytes()
        items.append((n.encode(), tfbundle._entry_proto(tensors[n].shape, len(data), len(raw), tfbundle.mask_crc(tfbundle.crc32c(raw)))))
        data += raw
    open(prefix + '.data-00000-of-00001', 'wb').write(bytes(data))
    blocks, handles, off = [], [], 0
    for part in (items[:10], items[10:25], items[25:]):
        b = tfbundle._with_trailer(tfbundle._build_block(part, restart_interval=4))
        handles.append((part[-1][0], tfbundle._varint(off) + tfbundle._varint(len(b) - 5)))
        blocks.append(b)
        off += len(b)
    meta = tfbundle._with_trailer(tfbundle._build_block([]))
    index = tfbundle._with_trailer(tfbundle._build_block(handles, restart_interval=1))
    footer = tfbundle._varint(off) + tfbundle._varint(len(meta) - 5) + tfbundle._varint(off + len(meta)) + tfbundle._varint(len(index) - 5)
    footer += b'\x00' * (40 - len(footer)) + struct.pack('<Q', tfbundle.MAGIC)
    open(prefix + '.index', 'wb').write(b''.join(blocks) + meta + index + footer)
    back = tfbundle.read_bundle(prefix)
    assert sorted(back) == sorted(tensors) and all(np.array_equal(back[n], tensors[n]) for n in tensors)
