/*
 * Reads the repeated Model.Metadata entries of a protobuf payload on the host, starting at the offset the
 * GPU decoder reports (metadata_span = offset of the first Metadata field | entry count << 32).  Metadata is
 * rare and small; the GPU has already validated it.
 * NOT COMPILED in this repository (no JDK in the build container).
 */
package com.sitewhere.gpu;

import java.nio.charset.StandardCharsets;
import java.util.HashMap;
import java.util.Map;

final class PbMetadata {

    static Map<String, String> read(byte[] p, long span) {
	Map<String, String> out = new HashMap<>();
	int pos = (int) (span & 0xffffffffL);
	int count = (int) (span >>> 32);
	final int metaTag = p[pos] & 0xff; // the first entry's tag identifies the metadata field of this message
	while (count > 0 && pos < p.length) {
	    int[] v = varint(p, pos);
	    int tag = v[0];
	    pos = v[1];
	    int wt = tag & 7;
	    if (tag == metaTag) {
		v = varint(p, pos);
		int len = v[0];
		pos = v[1];
		int end = pos + len;
		String name = null, value = null;
		while (pos < end) {
		    v = varint(p, pos);
		    int t2 = v[0];
		    pos = v[1];
		    v = varint(p, pos);
		    int l2 = v[0];
		    pos = v[1];
		    String s = new String(p, pos, l2, StandardCharsets.UTF_8);
		    pos += l2;
		    if (t2 == 10) {
			name = s;
		    } else if (t2 == 18) {
			value = s;
		    }
		}
		out.put(name, value);
		count--;
	    } else if (wt == 0) {
		pos = varint(p, pos)[1];
	    } else if (wt == 1) {
		pos += 8;
	    } else if (wt == 5) {
		pos += 4;
	    } else if (wt == 2) {
		v = varint(p, pos);
		pos = v[1] + v[0];
	    } else {
		break;
	    }
	}
	return out;
    }

    private static int[] varint(byte[] p, int pos) {
	long r = 0;
	int shift = 0;
	while (true) {
	    int b = p[pos++] & 0xff;
	    r |= (long) (b & 0x7f) << shift;
	    if ((b & 0x80) == 0) {
		break;
	    }
	    shift += 7;
	}
	return new int[] { (int) r, pos };
    }

    private PbMetadata() {
    }
}
