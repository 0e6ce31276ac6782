"""Oracle (test infrastructure, never on the product path): CPU restatement of the classification
postprocess of the reference example, /root/reference/src/python/examples/image_client.py:196-216.

Pinned against the reference's own function (stdout captured) by tests/test_oracle.py when
/root/reference is present, and against tests/golden/postprocess_golden.json everywhere.
"""
import numpy as np


def lines(output_array, batch_size, supports_batching):
    """The text postprocess() prints, one list entry per printed line; raises what it raises.

    :201-205  a batching model must return one row per request image
    :208-210  a non-batching model's output is one row
    :211-215  np.object_ elements are bytes: one character per byte (Latin-1), split on ':'
    :216      "    {score} ({index}) = {label}" from the first three fields
    """
    if supports_batching and len(output_array) != batch_size:
        raise Exception("expected {} results, got {}".format(batch_size, len(output_array)))
    out = []
    is_object = output_array.dtype.type == np.object_
    for row in output_array:
        row = row if supports_batching else [row]
        for element in row:
            text = "".join(map(chr, element)) if is_object else element
            fields = text.split(":")
            out.append("    {} ({}) = {}".format(fields[0], fields[1], fields[2]))
    return out
