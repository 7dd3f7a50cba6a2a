/*
 * sw_oracle_json.c — CPU oracle, part 2: JsonDeviceRequestDecoder.decode.
 * TEST INFRASTRUCTURE ONLY (see sw_oracle.h).  PARITY UNPINNED BY THE REFERENCE (see sw_oracle.h).
 *
 * Follows the reference's structure:  MAPPER.readValue(payload, DecodedDeviceRequest.class)
 *   (ES/decoder/json/JsonDeviceRequestDecoder.java:48-62)
 *   -> JsonDeviceRequestMarshaler.deserialize: readTree, then get("type"/"request"/"deviceToken"/"originator")
 *      (ES/decoder/json/JsonDeviceRequestMarshaler.java:57-84)
 *   -> unmarshal: MAPPER.treeToValue(json, XxxCreateRequest.class) with a default ObjectMapper (:97-160)
 * over the POJOs sitewhere-core-api/.../rest/model/device/event/request/Device{Event,Measurement,
 * Location,Alert}CreateRequest.java.  Jackson 2.9.x behaviour (third-party, recalled) is restated:
 * a strict JSON tokenizer (UTF8StreamJsonParser defaults), a node tree with last-wins duplicate keys
 * (JsonNodeDeserializer), bean deserialization with FAIL_ON_UNKNOWN_PROPERTIES and scalar coercions
 * (StdDeserializer._parseXxx, EnumDeserializer, StringDeserializer, DateDeserializer + StdDateFormat,
 * MapDeserializer), NumberInput.parseDouble.  Double.parseDouble is glibc strtod (both are correctly
 * rounded, round-half-even).
 *
 * Records the GPU path deliberately leaves to the stock decoder get ST_PASSTHROUGH (same rules on
 * both sides, listed in DESIGN.md): non-UTF-8 encodings, nesting deeper than 256, non-event request
 * types, escaped keys at depth <= 2, key-forming strings (deviceToken, measurement name, alert type)
 * with escapes or coerced from non-strings, numeric / enum / boolean / date strings with escapes,
 * numeric strings outside the plain decimal grammar, date strings that would go to the RFC-1123
 * parser, years before 1583.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "sw_oracle.h"

#define ST_OK 0
#define ST_DECODE 1
#define ST_RUNTIME 2
#define ST_PASS 3

#define KIND_MEASUREMENT 0u
#define KIND_LOCATION 1u
#define KIND_ALERT 2u
#define F_UPDATE_STATE (1u << 0)
#define F_HAS_DATE (1u << 1)
#define F_HAS_ORIGINATOR (1u << 2)
#define F_HAS_METADATA (1u << 3)
#define F_HAS_ALT_ID (1u << 4)
#define F_NULL_TOKEN (1u << 5)
#define F_HAS_V0 (1u << 6)
#define F_HAS_V1 (1u << 7)
#define F_HAS_V2 (1u << 8)
#define F_STR_ESCAPED (1u << 9)
#define F_HAS_NAME (1u << 10)
#define F_HAS_MESSAGE (1u << 11)
#define F_NAME_COERCED (1u << 12)
#define F_NULL_REQUEST (1u << 13)
#define F_NULL_LEVEL (1u << 14)
#define F_NULL_SOURCE (1u << 15)

#define MAX_DEPTH 256

/* ---------------------------------------------------------------- node tree */
enum { N_OBJECT, N_ARRAY, N_STRING, N_INT, N_FLOAT, N_TRUE, N_FALSE, N_NULL };
typedef struct jnode {
    int type;
    uint32_t off, len; /* STRING: the bytes between the quotes; numbers/literals: the token; containers: whole text */
    int escaped;       /* STRING: contains a backslash */
    int nonascii;      /* STRING: contains a byte >= 0x80 */
    struct jfield *fields; /* OBJECT */
    size_t nfields;
} jnode;
typedef struct jfield {
    uint32_t koff, klen;
    int kescaped;
    jnode *value;
} jfield;

typedef struct {
    const uint8_t *s;
    uint32_t n, pos;
    int err;      /* JsonParseException */
    int too_deep; /* depth > MAX_DEPTH */
    int esc_key;  /* an escaped key at depth <= 2 */
    /* arena */
    jnode **nodes;
    size_t nnodes, capnodes;
} jp;

static jnode *jnew(jp *p, int type) {
    jnode *n = (jnode *)calloc(1, sizeof *n);
    n->type = type;
    if (p->nnodes == p->capnodes) {
        p->capnodes = p->capnodes ? p->capnodes * 2 : 32;
        p->nodes = (jnode **)realloc(p->nodes, p->capnodes * sizeof(jnode *));
    }
    p->nodes[p->nnodes++] = n;
    return n;
}
static void jfree_all(jp *p) {
    for (size_t i = 0; i < p->nnodes; i++) {
        free(p->nodes[i]->fields);
        free(p->nodes[i]);
    }
    free(p->nodes);
}
static void skip_ws(jp *p) {
    while (p->pos < p->n) {
        uint8_t c = p->s[p->pos];
        if (c == ' ' || c == '\t' || c == '\n' || c == '\r')
            p->pos++;
        else
            break;
    }
}
static int is_hex(uint8_t c) { return (c >= '0' && c <= '9') || (c >= 'a' && c <= 'f') || (c >= 'A' && c <= 'F'); }

/* scans a string body starting after the opening quote; returns 0 on error */
static int scan_string(jp *p, uint32_t *off, uint32_t *len, int *escaped, int *nonascii) {
    uint32_t start = p->pos;
    *escaped = 0;
    *nonascii = 0;
    for (;;) {
        if (p->pos >= p->n) return 0; /* unexpected end-of-input in string */
        uint8_t c = p->s[p->pos];
        if (c == '"') break;
        if (c < 0x20) return 0; /* illegal unquoted control character */
        if (c == '\\') {
            *escaped = 1;
            p->pos++;
            if (p->pos >= p->n) return 0;
            uint8_t e = p->s[p->pos];
            if (e == 'u') {
                if (p->pos + 4 >= p->n) return 0;
                for (int k = 1; k <= 4; k++)
                    if (!is_hex(p->s[p->pos + k])) return 0;
                p->pos += 5;
            } else if (e == '"' || e == '\\' || e == '/' || e == 'b' || e == 'f' || e == 'n' || e == 'r' || e == 't') {
                p->pos++;
            } else
                return 0; /* unrecognized character escape */
            continue;
        }
        if (c < 0x80) {
            p->pos++;
            continue;
        }
        /* UTF-8: Jackson checks the lead byte class and that continuation bytes are 10xxxxxx */
        *nonascii = 1;
        int need;
        if ((c & 0xE0) == 0xC0)
            need = 1;
        else if ((c & 0xF0) == 0xE0)
            need = 2;
        else if ((c & 0xF8) == 0xF0)
            need = 3;
        else
            return 0; /* invalid UTF-8 start byte */
        p->pos++;
        for (int k = 0; k < need; k++) {
            if (p->pos >= p->n) return 0;
            if ((p->s[p->pos] & 0xC0) != 0x80) return 0;
            p->pos++;
        }
    }
    *off = start;
    *len = p->pos - start;
    p->pos++; /* closing quote */
    return 1;
}

/* Character.isJavaIdentifierPart for the bytes that can follow a literal / number (ASCII part;
 * any byte >= 0x80 starts a non-ASCII char, which Jackson reports as an error here either way) */
static int ident_part(uint8_t c) {
    return (c >= '0' && c <= '9') || (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || c == '_' || c == '$' ||
           c >= 0x80 || c < 0x09 || (c >= 0x0e && c <= 0x1b) || c == 0x7f;
}

static jnode *parse_value(jp *p, int depth);

static jnode *parse_number(jp *p) {
    uint32_t start = p->pos;
    int is_float = 0;
    if (p->s[p->pos] == '-') {
        p->pos++;
        if (p->pos >= p->n) goto bad;
    }
    if (p->s[p->pos] == '0') {
        p->pos++;
        if (p->pos < p->n && p->s[p->pos] >= '0' && p->s[p->pos] <= '9') goto bad; /* leading zeroes not allowed */
    } else if (p->s[p->pos] >= '1' && p->s[p->pos] <= '9') {
        while (p->pos < p->n && p->s[p->pos] >= '0' && p->s[p->pos] <= '9') p->pos++;
    } else
        goto bad;
    if (p->pos < p->n && p->s[p->pos] == '.') {
        is_float = 1;
        p->pos++;
        if (p->pos >= p->n || p->s[p->pos] < '0' || p->s[p->pos] > '9') goto bad; /* decimal point not followed by a digit */
        while (p->pos < p->n && p->s[p->pos] >= '0' && p->s[p->pos] <= '9') p->pos++;
    }
    if (p->pos < p->n && (p->s[p->pos] == 'e' || p->s[p->pos] == 'E')) {
        is_float = 1;
        p->pos++;
        if (p->pos < p->n && (p->s[p->pos] == '+' || p->s[p->pos] == '-')) p->pos++;
        if (p->pos >= p->n || p->s[p->pos] < '0' || p->s[p->pos] > '9') goto bad; /* exponent has no digits */
        while (p->pos < p->n && p->s[p->pos] >= '0' && p->s[p->pos] <= '9') p->pos++;
    }
    {
        jnode *n = jnew(p, is_float ? N_FLOAT : N_INT);
        n->off = start;
        n->len = p->pos - start;
        return n;
    }
bad:
    p->err = 1;
    return NULL;
}

static jnode *parse_literal(jp *p, const char *lit, int type) {
    size_t l = strlen(lit);
    if (p->pos + l > p->n || memcmp(p->s + p->pos, lit, l) != 0) {
        p->err = 1;
        return NULL;
    }
    uint32_t start = p->pos;
    p->pos += (uint32_t)l;
    if (p->pos < p->n && ident_part(p->s[p->pos])) { /* _checkMatchEnd: "Unrecognized token" */
        p->err = 1;
        return NULL;
    }
    jnode *n = jnew(p, type);
    n->off = start;
    n->len = (uint32_t)l;
    return n;
}

static jnode *parse_value(jp *p, int depth) {
    skip_ws(p);
    if (p->pos >= p->n) {
        p->err = 1;
        return NULL;
    }
    uint8_t c = p->s[p->pos];
    if (c == '{' || c == '[') {
        if (depth + 1 > MAX_DEPTH) {
            p->too_deep = 1;
            return NULL;
        }
        uint32_t start = p->pos;
        p->pos++;
        jnode *n = jnew(p, c == '{' ? N_OBJECT : N_ARRAY);
        size_t cap = 0;
        skip_ws(p);
        uint8_t close = c == '{' ? '}' : ']';
        if (p->pos < p->n && p->s[p->pos] == close) {
            p->pos++;
        } else {
            for (;;) {
                uint32_t koff = 0, klen = 0;
                int kesc = 0, knon = 0;
                if (c == '{') {
                    skip_ws(p);
                    if (p->pos >= p->n || p->s[p->pos] != '"') { /* was expecting double-quote to start field name */
                        p->err = 1;
                        return NULL;
                    }
                    p->pos++;
                    if (!scan_string(p, &koff, &klen, &kesc, &knon)) {
                        p->err = 1;
                        return NULL;
                    }
                    if (kesc && depth + 1 <= 2) p->esc_key = 1;
                    skip_ws(p);
                    if (p->pos >= p->n || p->s[p->pos] != ':') {
                        p->err = 1;
                        return NULL;
                    }
                    p->pos++;
                }
                jnode *v = parse_value(p, depth + 1);
                if (!v) return NULL;
                if (c == '{') {
                    /* JsonNodeDeserializer: a duplicate field replaces the earlier value */
                    size_t k;
                    for (k = 0; k < n->nfields; k++)
                        if (n->fields[k].klen == klen && memcmp(p->s + n->fields[k].koff, p->s + koff, klen) == 0) break;
                    if (k == n->nfields) {
                        if (n->nfields == cap) {
                            cap = cap ? cap * 2 : 8;
                            n->fields = (jfield *)realloc(n->fields, cap * sizeof(jfield));
                        }
                        n->nfields++;
                    }
                    n->fields[k].koff = koff;
                    n->fields[k].klen = klen;
                    n->fields[k].kescaped = kesc;
                    n->fields[k].value = v;
                }
                skip_ws(p);
                if (p->pos >= p->n) {
                    p->err = 1;
                    return NULL;
                }
                if (p->s[p->pos] == ',') {
                    p->pos++;
                    if (c == '[') {
                        skip_ws(p);
                        if (p->pos < p->n && p->s[p->pos] == ']') { /* trailing comma */
                            p->err = 1;
                            return NULL;
                        }
                    }
                    continue;
                }
                if (p->s[p->pos] == close) {
                    p->pos++;
                    break;
                }
                p->err = 1;
                return NULL;
            }
        }
        n->off = start;
        n->len = p->pos - start;
        return n;
    }
    if (c == '"') {
        p->pos++;
        jnode *n = jnew(p, N_STRING);
        if (!scan_string(p, &n->off, &n->len, &n->escaped, &n->nonascii)) {
            p->err = 1;
            return NULL;
        }
        return n;
    }
    if (c == '-' || (c >= '0' && c <= '9')) {
        jnode *n = parse_number(p);
        if (!n) return NULL;
        /* a number must be followed by a separator: anything else is reported when the next token
         * is read; at root level _verifyRootSpace demands whitespace */
        if (p->pos < p->n) {
            uint8_t d = p->s[p->pos];
            if (depth == 0) {
                if (!(d == ' ' || d == '\t' || d == '\n' || d == '\r')) {
                    p->err = 1;
                    return NULL;
                }
            }
        }
        return n;
    }
    if (c == 't') return parse_literal(p, "true", N_TRUE);
    if (c == 'f') return parse_literal(p, "false", N_FALSE);
    if (c == 'n') return parse_literal(p, "null", N_NULL);
    p->err = 1;
    return NULL;
}

static jnode *obj_get(const uint8_t *s, jnode *o, const char *key) {
    size_t kl = strlen(key);
    for (size_t i = 0; i < o->nfields; i++)
        if (o->fields[i].klen == kl && memcmp(s + o->fields[i].koff, key, kl) == 0) return o->fields[i].value;
    return NULL;
}
static int str_eq(const uint8_t *s, const jnode *n, const char *lit) {
    size_t l = strlen(lit);
    return n->len == l && memcmp(s + n->off, lit, l) == 0;
}

/* ---------------------------------------------------------------- numbers */
/* integer token -> low 64 bits (two's complement), "is zero", and the correctly rounded double */
static void int_token(const uint8_t *s, uint32_t len, uint64_t *low64, double *dv) {
    int neg = s[0] == '-';
    uint64_t v = 0;
    for (uint32_t i = neg; i < len; i++) v = v * 10 + (uint64_t)(s[i] - '0');
    *low64 = neg ? (uint64_t)0 - v : v;
    char buf[64];
    char *b = len < sizeof buf ? buf : (char *)malloc(len + 1);
    memcpy(b, s, len);
    b[len] = 0;
    double d = strtod(b, NULL); /* IntNode/LongNode/BigIntegerNode.doubleValue(): nearest double */
    if (d == 0) d = 0.0;        /* "-0" is the int 0: +0.0 */
    *dv = d;
    if (b != buf) free(b);
}
int orc_json_number_to_double(const char *tok, uint32_t len, double *out) {
    /* NumberInput.parseDouble: the one special-cased literal, else Double.parseDouble */
    static const char nasty[] = "2.2250738585072012e-308";
    int is_float = 0;
    for (uint32_t i = 0; i < len; i++)
        if (tok[i] == '.' || tok[i] == 'e' || tok[i] == 'E') is_float = 1;
    if (!is_float) {
        uint64_t lo;
        int_token((const uint8_t *)tok, len, &lo, out);
        return 0;
    }
    if (len == sizeof nasty - 1 && memcmp(tok, nasty, len) == 0) {
        uint64_t one = 1;
        memcpy(out, &one, 8); /* Double.MIN_VALUE */
        return 0;
    }
    char buf[64];
    char *b = len < sizeof buf ? buf : (char *)malloc(len + 1);
    memcpy(b, tok, len);
    b[len] = 0;
    *out = strtod(b, NULL);
    if (b != buf) free(b);
    return 0;
}

/* Java String.trim(): strip chars <= 0x20 at both ends */
static void jtrim(const uint8_t *s, uint32_t *off, uint32_t *len) {
    while (*len && s[*off] <= 0x20) {
        (*off)++;
        (*len)--;
    }
    while (*len && s[*off + *len - 1] <= 0x20) (*len)--;
}

/* string -> double for numeric fields.  returns 0 ok, 1 decode error, 2 pass-through.
 * *isnull is set for "" / "null" (primitive: 0.0, wrapper: null) */
static int string_to_double(const uint8_t *s, const jnode *n, double *out, int *isnull) {
    *isnull = 0;
    if (n->escaped) return 2;
    uint32_t off = n->off, len = n->len;
    jtrim(s, &off, &len);
    const uint8_t *t = s + off;
    if (len == 0 || (len == 4 && memcmp(t, "null", 4) == 0)) {
        *isnull = 1;
        *out = 0.0;
        return 0;
    }
    if ((len == 8 && memcmp(t, "Infinity", 8) == 0) || (len == 3 && memcmp(t, "INF", 3) == 0)) {
        *out = INFINITY;
        return 0;
    }
    if ((len == 9 && memcmp(t, "-Infinity", 9) == 0) || (len == 4 && memcmp(t, "-INF", 4) == 0)) {
        *out = -INFINITY;
        return 0;
    }
    if (len == 3 && memcmp(t, "NaN", 3) == 0) {
        uint64_t b = 0x7ff8000000000000ull;
        memcpy(out, &b, 8);
        return 0;
    }
    /* plain decimal grammar of Double.parseDouble: [+-]? (digits [. digits?]? | . digits) ([eE][+-]?digits)? */
    uint32_t i = 0;
    if (t[i] == '+' || t[i] == '-') i++;
    uint32_t nd = 0;
    while (i < len && t[i] >= '0' && t[i] <= '9') {
        i++;
        nd++;
    }
    if (i < len && t[i] == '.') {
        i++;
        while (i < len && t[i] >= '0' && t[i] <= '9') {
            i++;
            nd++;
        }
    }
    if (nd == 0) return 2; /* not the plain grammar (could still be hex / garbage): stock decoder decides */
    if (i < len && (t[i] == 'e' || t[i] == 'E')) {
        i++;
        if (i < len && (t[i] == '+' || t[i] == '-')) i++;
        uint32_t ne = 0;
        while (i < len && t[i] >= '0' && t[i] <= '9') {
            i++;
            ne++;
        }
        if (ne == 0) return 2;
    }
    if (i != len) return 2;
    char buf[64];
    char *b = len < sizeof buf ? buf : (char *)malloc(len + 1);
    memcpy(b, t, len);
    b[len] = 0;
    *out = strtod(b, NULL);
    if (b != buf) free(b);
    return 0;
}

/* a numeric property (double / Double).  rc: 0 ok, 1 decode error, 2 pass-through */
static int node_to_double(const uint8_t *s, const jnode *n, double *out, int *isnull) {
    *isnull = 0;
    switch (n->type) {
    case N_INT:
    case N_FLOAT: orc_json_number_to_double((const char *)s + n->off, n->len, out); return 0;
    case N_NULL:
        *isnull = 1;
        *out = 0.0;
        return 0;
    case N_STRING: return string_to_double(s, n, out, isnull);
    default: return 1; /* booleans, arrays, objects: handleUnexpectedToken */
    }
}

/* ---------------------------------------------------------------- dates */
static int64_t days_from_civil(int64_t y, int64_t m, int64_t d) {
    y -= m <= 2;
    int64_t era = (y >= 0 ? y : y - 399) / 400;
    int64_t yoe = y - era * 400;
    int64_t doy = (153 * (m + (m > 2 ? -3 : 9)) + 2) / 5 + d - 1;
    int64_t doe = yoe * 365 + yoe / 4 - yoe / 100 + doy;
    return era * 146097 + doe - 719468;
}
static int d2(const char *s) { return (s[0] - '0') * 10 + (s[1] - '0'); }
static int isd(char c) { return c >= '0' && c <= '9'; }
/* lenient GregorianCalendar.set(year, month0, day, h, m, s) in UTC, year >= 1583 */
static int64_t lenient_utc_ms(int year, int month0, int day, int hour, int min, int sec, int ms) {
    int64_t y = year + (month0 >= 0 ? month0 / 12 : -((11 - month0) / 12));
    int64_t m = ((month0 % 12) + 12) % 12;
    int64_t days = days_from_civil(y, m + 1, 1) + (day - 1);
    return ((days * 24 + hour) * 60 + min) * 60000 + (int64_t)sec * 1000 + ms;
}
/* StdDateFormat (Jackson 2.9.5) on an ASCII string without escapes.
 * rc: 0 ok, 1 error (IllegalArgumentException -> decode failure), 2 pass-through */
int orc_parse_date(const char *s, uint32_t len, int64_t *out_ms) {
    /* looksLikeISO8601 */
    if (len >= 7 && isd(s[0]) && isd(s[3]) && s[4] == '-' && isd(s[5])) {
        if (len <= 10) {
            /* PATTERN_PLAIN \d\d\d\d-\d\d-\d\d */
            if (len == 10 && isd(s[0]) && isd(s[1]) && isd(s[2]) && isd(s[3]) && s[4] == '-' && isd(s[5]) && isd(s[6]) &&
                s[7] == '-' && isd(s[8]) && isd(s[9])) {
                int year = d2(s) * 100 + d2(s + 2);
                if (year < 1583) return 2;
                *out_ms = lenient_utc_ms(year, d2(s + 5) - 1, d2(s + 8), 0, 0, 0, 0);
                return 0;
            }
            return 1;
        }
        /* PATTERN_ISO8601: \d{4}-\d\d-\d\dT\d\d:\d\d(:\d\d)?(\.\d+)?(Z|[+-]\d\d(:?\d\d)?)? */
        uint32_t i = 0;
        if (!(len >= 16 && isd(s[0]) && isd(s[1]) && isd(s[2]) && isd(s[3]) && s[4] == '-' && isd(s[5]) && isd(s[6]) &&
              s[7] == '-' && isd(s[8]) && isd(s[9]) && s[10] == 'T' && isd(s[11]) && isd(s[12]) && s[13] == ':' &&
              isd(s[14]) && isd(s[15])))
            return 1;
        i = 16;
        int seconds = 0;
        if (i < len && s[i] == ':') {
            if (i + 3 > len || !isd(s[i + 1]) || !isd(s[i + 2])) return 1;
            seconds = d2(s + i + 1);
            i += 3;
        }
        int msecs = 0;
        if (i < len && s[i] == '.') {
            uint32_t j = i + 1;
            while (j < len && isd(s[j])) j++;
            uint32_t fl = j - (i + 1);
            if (fl == 0) return 1; /* \.\d+ needs a digit; then the rest cannot match */
            if (fl > 9) return 1;  /* "only allow up to nanos" */
            if (fl >= 1) msecs += 100 * (s[i + 1] - '0');
            if (fl >= 2) msecs += 10 * (s[i + 2] - '0');
            if (fl >= 3) msecs += (s[i + 3] - '0');
            i = j;
        }
        int64_t offset_ms = 0;
        if (i < len) {
            if (s[i] == 'Z') {
                i++;
            } else if (s[i] == '+' || s[i] == '-') {
                uint32_t st = i;
                if (i + 3 > len || !isd(s[i + 1]) || !isd(s[i + 2])) return 1;
                int hh = d2(s + i + 1), mm = 0;
                i += 3;
                if (i < len) {
                    if (s[i] == ':') i++;
                    if (i + 2 > len || !isd(s[i]) || !isd(s[i + 1])) return 1;
                    mm = d2(s + i);
                    i += 2;
                }
                offset_ms = ((int64_t)hh * 3600 + (int64_t)mm * 60) * 1000;
                if (s[st] == '-') offset_ms = -offset_ms;
            } else
                return 1;
        }
        if (i != len) return 1;
        int year = d2(s) * 100 + d2(s + 2);
        if (year < 1583) return 2;
        *out_ms = lenient_utc_ms(year, d2(s + 5) - 1, d2(s + 8), d2(s + 11), d2(s + 14), seconds, msecs) - offset_ms;
        return 0;
    }
    /* "stringified" timestamp: all digits, optionally a leading '-' */
    {
        int all = len > 0;
        for (uint32_t k = 0; k < len; k++) {
            if (isd(s[k])) continue;
            if (k == 0 && s[k] == '-') continue;
            all = 0;
            break;
        }
        if (all) {
            int neg = s[0] == '-';
            uint32_t nd = len - neg;
            if (neg) {
                /* Long.parseLong of a negative string */
                if (nd == 0 || nd > 19) return 1;
            } else {
                /* NumberInput.inLongRange(str, false) else -> RFC 1123 parser */
                if (nd > 19) return 2;
            }
            unsigned __int128 v = 0;
            for (uint32_t k = neg; k < len; k++) v = v * 10 + (unsigned)(s[k] - '0');
            if (neg) {
                if (v > (unsigned __int128)9223372036854775808ull) return 1;
                *out_ms = (int64_t)(0 - (uint64_t)v);
            } else {
                if (v > (unsigned __int128)9223372036854775807ull) return 2; /* not in long range -> RFC 1123 */
                *out_ms = (int64_t)(uint64_t)v;
            }
            return 0;
        }
    }
    return 2; /* RFC-1123 parser territory */
}

/* ---------------------------------------------------------------- enums / booleans */
static const char *E_EVENT_TYPE[] = {"Measurement", "Location", "Alert", "CommandInvocation", "CommandResponse", "StateChange"};
static const char *E_SOURCE[] = {"Device", "System"};
static const char *E_LEVEL[] = {"Info", "Warning", "Error", "Critical"};

/* rc 0 ok (*isnull for JSON null), 1 error, 2 pass-through */
static int node_to_enum(const uint8_t *s, const jnode *n, const char **names, int count, int *ordinal, int *isnull) {
    *isnull = 0;
    if (n->type == N_NULL) {
        *isnull = 1;
        return 0;
    }
    if (n->type == N_STRING) {
        if (n->escaped) return 2;
        for (int k = 0; k < count; k++)
            if (str_eq(s, n, names[k])) {
                *ordinal = k;
                return 0;
            }
        /* _deserializeAltString: trimmed; a string index "2" is accepted */
        uint32_t off = n->off, len = n->len;
        jtrim(s, &off, &len);
        if (len > 0 && isd((char)s[off])) {
            uint64_t v = 0;
            int ok = 1;
            for (uint32_t k = 0; k < len; k++) {
                if (!isd((char)s[off + k])) {
                    ok = 0;
                    break;
                }
                v = v * 10 + (uint64_t)(s[off + k] - '0');
                if (v > 2147483647ull) {
                    ok = 0; /* Integer.parseInt overflow -> NumberFormatException, ignored */
                    break;
                }
            }
            if (ok && v < (uint64_t)count) {
                *ordinal = (int)v;
                return 0;
            }
        }
        return 1;
    }
    if (n->type == N_INT) {
        /* p.getIntValue() of the tree parser: the node's intValue(), i.e. the low 32 bits */
        uint64_t lo;
        double dv;
        int_token(s + n->off, n->len, &lo, &dv);
        int32_t idx = (int32_t)(uint32_t)lo;
        if (idx >= 0 && idx < count) {
            *ordinal = idx;
            return 0;
        }
        return 1;
    }
    return 1;
}

static int node_to_bool(const uint8_t *s, const jnode *n, int *val) {
    switch (n->type) {
    case N_TRUE: *val = 1; return 0;
    case N_FALSE: *val = 0; return 0;
    case N_NULL: *val = 0; return 0;
    case N_INT: {
        /* !"0".equals(p.getText()): the node's canonical text */
        uint64_t lo;
        double dv;
        int_token(s + n->off, n->len, &lo, &dv);
        int zero = 1;
        for (uint32_t k = 0; k < n->len; k++)
            if (s[n->off + k] >= '1' && s[n->off + k] <= '9') zero = 0;
        *val = !zero;
        return 0;
    }
    case N_STRING: {
        if (n->escaped) return 2;
        uint32_t off = n->off, len = n->len;
        jtrim(s, &off, &len);
        const uint8_t *t = s + off;
        if (len == 4 && (memcmp(t, "true", 4) == 0 || memcmp(t, "True", 4) == 0)) {
            *val = 1;
            return 0;
        }
        if (len == 5 && (memcmp(t, "false", 5) == 0 || memcmp(t, "False", 5) == 0)) {
            *val = 0;
            return 0;
        }
        if (len == 0 || (len == 4 && memcmp(t, "null", 4) == 0)) {
            *val = 0;
            return 0;
        }
        return 1;
    }
    default: return 1;
    }
}

/* String property: rc 0 ok, 1 error.  *present = 0 for null. *coerced for scalar non-strings */
static int node_to_string(const jnode *n, int *present, int *coerced, int *escaped) {
    *present = 0;
    *coerced = 0;
    *escaped = 0;
    switch (n->type) {
    case N_NULL: return 0;
    case N_STRING:
        *present = 1;
        *escaped = n->escaped;
        return 0;
    case N_INT:
    case N_FLOAT:
    case N_TRUE:
    case N_FALSE:
        *present = 1;
        *coerced = 1;
        return 0;
    default: return 1;
    }
}

static uint64_t mk_meta(uint32_t kind, uint32_t flags, uint32_t level, uint32_t source, int has_etype, uint32_t etype,
                        uint32_t rec) {
    return (uint64_t)(kind & 0xff) | ((uint64_t)(flags & 0xffff) << 8) | ((uint64_t)(level & 3) << 24) |
           ((uint64_t)(source & 1) << 26) | ((uint64_t)(has_etype ? 1 : 0) << 27) | ((uint64_t)(etype & 7) << 28) |
           ((uint64_t)rec << 32);
}
static uint64_t mk_span(uint64_t off, uint64_t len) { return (off & 0xffffffffu) | (len << 32); }

typedef struct {
    uint64_t meta, th, nh, tspan, nspan, ospan, mspan, mdspan, aspan;
    int64_t date;
    double v0, v1, v2;
} jrow;

static const char *BASE_PROPS[] = {"alternateId", "eventType", "eventDate", "updateState", "metadata"};
static int in_list(const uint8_t *s, const jfield *f, const char **list, int n) {
    for (int k = 0; k < n; k++) {
        size_t l = strlen(list[k]);
        if (f->klen == l && memcmp(s + f->koff, list[k], l) == 0) return 1;
    }
    return 0;
}

/* one record.  Returns status; fills *row when ST_OK */
static int decode_one(const uint8_t *s, uint32_t n, uint32_t rec, jrow *row) {
    memset(row, 0, sizeof *row);
    /* ByteSourceJsonBootstrapper.detectEncoding: UTF-16/32 by BOM or zero bytes */
    uint32_t start = 0;
    if (n >= 2) {
        if (s[0] == 0 || s[1] == 0) return ST_PASS;
        if ((s[0] == 0xFE && s[1] == 0xFF) || (s[0] == 0xFF && s[1] == 0xFE)) return ST_PASS;
    }
    if (n >= 3 && s[0] == 0xEF && s[1] == 0xBB && s[2] == 0xBF) start = 3;
    jp p;
    memset(&p, 0, sizeof p);
    p.s = s;
    p.n = n;
    p.pos = start;
    int status = ST_DECODE;
    jnode *root = NULL;
    skip_ws(&p);
    if (p.pos >= p.n) goto done; /* "No content to map due to end-of-input" */
    root = parse_value(&p, 0);
    if (!root) {
        status = p.too_deep ? ST_PASS : ST_DECODE;
        goto done;
    }
    if (root->type != N_OBJECT) goto done; /* node.get("type") == null -> "Event type is required." */
    if (p.esc_key) {
        status = ST_PASS;
        goto done;
    }
    {
        jnode *typeNode = obj_get(s, root, "type");
        if (!typeNode) goto done;
        if (typeNode->type != N_STRING) goto done; /* asText() of anything else is not an enum constant */
        if (typeNode->escaped) {
            status = ST_PASS;
            goto done;
        }
        static const char *TYPES[] = {"RegisterDevice", "DeviceLocation", "DeviceAlert", "DeviceMeasurement",
                                      "DeviceStream",   "DeviceStreamData", "Acknowledge"};
        int ty = -1;
        for (int k = 0; k < 7; k++)
            if (str_eq(s, typeNode, TYPES[k])) ty = k;
        if (ty < 0) goto done; /* "Event type is not valid." */
        jnode *request = obj_get(s, root, "request");
        if (!request) goto done; /* "Request is missing." */
        jnode *tokenNode = obj_get(s, root, "deviceToken");
        if (!tokenNode) goto done; /* "Device token is missing." */
        jnode *origNode = obj_get(s, root, "originator");
        if (ty != 1 && ty != 2 && ty != 3) {
            status = ST_PASS;
            goto done;
        }
        uint32_t kind = ty == 1 ? KIND_LOCATION : (ty == 2 ? KIND_ALERT : KIND_MEASUREMENT);
        uint32_t flags = 0;
        int pass = 0;
        /* deviceToken.textValue(): null unless textual */
        if (tokenNode->type == N_STRING) {
            if (tokenNode->escaped) pass = 1;
            row->tspan = mk_span(tokenNode->off, tokenNode->len);
            row->th = orc_hash64(s + tokenNode->off, tokenNode->len);
        } else
            flags |= F_NULL_TOKEN;
        if (origNode && origNode->type == N_STRING) {
            flags |= F_HAS_ORIGINATOR;
            if (origNode->escaped) flags |= F_STR_ESCAPED;
            row->ospan = mk_span(origNode->off, origNode->len);
        }
        if (request->type == N_NULL) {
            /* treeToValue(NullNode) == null: a DecodedDeviceRequest with a null request */
            if (pass) {
                status = ST_PASS;
                goto done;
            }
            row->meta = mk_meta(kind, flags | F_NULL_REQUEST, 0, 0, 0, 0, rec);
            status = ST_OK;
            goto done;
        }
        if (request->type != N_OBJECT) goto done; /* no creator from scalar / array */
        /* ---- bean deserialization: defaults of the POJO constructors */
        int has_etype = 1;
        uint32_t etype = kind, level = 0, source = 0;
        static const char *MX_PROPS[] = {"name", "value"};
        static const char *LOC_PROPS[] = {"latitude", "longitude", "elevation"};
        static const char *AL_PROPS[] = {"source", "level", "type", "message"};
        const char **own = kind == KIND_MEASUREMENT ? MX_PROPS : (kind == KIND_LOCATION ? LOC_PROPS : AL_PROPS);
        int nown = kind == KIND_MEASUREMENT ? 2 : (kind == KIND_LOCATION ? 3 : 4);
        int err = 0;
        for (size_t k = 0; k < request->nfields && !err; k++) {
            jfield *f = &request->fields[k];
            if (!in_list(s, f, BASE_PROPS, 5) && !in_list(s, f, own, nown)) err = 1; /* UnrecognizedPropertyException */
        }
        for (size_t k = 0; k < request->nfields && !err; k++) {
            jfield *f = &request->fields[k];
            jnode *v = f->value;
            const uint8_t *key = s + f->koff;
            uint32_t kl = f->klen;
#define KEY(lit) (kl == sizeof(lit) - 1 && memcmp(key, lit, kl) == 0)
            if (KEY("alternateId")) {
                int present, coerced, escaped;
                if (node_to_string(v, &present, &coerced, &escaped)) {
                    err = 1;
                    break;
                }
                if (present) {
                    flags |= F_HAS_ALT_ID;
                    if (coerced) flags |= F_NAME_COERCED;
                    if (escaped) flags |= F_STR_ESCAPED;
                    row->aspan = mk_span(v->off, v->len);
                }
            } else if (KEY("eventType")) {
                int ord = 0, isnull;
                int rc = node_to_enum(s, v, E_EVENT_TYPE, 6, &ord, &isnull);
                if (rc == 1) err = 1;
                if (rc == 2) pass = 1;
                if (rc == 0) {
                    has_etype = !isnull;
                    etype = isnull ? 0 : (uint32_t)ord;
                }
            } else if (KEY("eventDate")) {
                if (v->type == N_NULL) {
                    /* null date */
                } else if (v->type == N_INT) {
                    uint64_t lo;
                    double dv;
                    int_token(s + v->off, v->len, &lo, &dv);
                    row->date = (int64_t)lo; /* getLongValue(): low 64 bits */
                    flags |= F_HAS_DATE;
                } else if (v->type == N_STRING) {
                    if (v->escaped || v->nonascii)
                        pass = 1;
                    else {
                        uint32_t off = v->off, len = v->len;
                        jtrim(s, &off, &len);
                        if (len == 0 || (len == 4 && memcmp(s + off, "null", 4) == 0)) {
                            /* empty / "null" -> null date */
                        } else {
                            int64_t ms;
                            int rc = orc_parse_date((const char *)s + off, len, &ms);
                            if (rc == 1) err = 1;
                            if (rc == 2) pass = 1;
                            if (rc == 0) {
                                row->date = ms;
                                flags |= F_HAS_DATE;
                            }
                        }
                    }
                } else
                    err = 1;
            } else if (KEY("updateState")) {
                int b = 0;
                int rc = node_to_bool(s, v, &b);
                if (rc == 1) err = 1;
                if (rc == 2) pass = 1;
                if (rc == 0 && b) flags |= F_UPDATE_STATE;
            } else if (KEY("metadata")) {
                if (v->type == N_NULL) {
                } else if (v->type == N_OBJECT) {
                    for (size_t m = 0; m < v->nfields; m++) {
                        int t = v->fields[m].value->type;
                        if (t == N_OBJECT || t == N_ARRAY) err = 1; /* String value from a container */
                    }
                    flags |= F_HAS_METADATA;
                    row->mdspan = mk_span(v->off, v->len);
                } else
                    err = 1;
            } else if (KEY("value") || KEY("latitude") || KEY("longitude") || KEY("elevation")) {
                double d;
                int isnull;
                int rc = node_to_double(s, v, &d, &isnull);
                if (rc == 1) err = 1;
                if (rc == 2) pass = 1;
                if (rc == 0) {
                    if (KEY("value")) {
                        row->v0 = d; /* primitive double: null -> 0.0 */
                    } else if (KEY("latitude")) {
                        row->v0 = isnull ? 0.0 : d;
                        flags = isnull ? (flags & ~F_HAS_V0) : (flags | F_HAS_V0);
                    } else if (KEY("longitude")) {
                        row->v1 = isnull ? 0.0 : d;
                        flags = isnull ? (flags & ~F_HAS_V1) : (flags | F_HAS_V1);
                    } else {
                        row->v2 = isnull ? 0.0 : d;
                        flags = isnull ? (flags & ~F_HAS_V2) : (flags | F_HAS_V2);
                    }
                }
            } else if (KEY("name") || (kind == KIND_ALERT && KEY("type"))) {
                int present, coerced, escaped;
                if (node_to_string(v, &present, &coerced, &escaped)) {
                    err = 1;
                    break;
                }
                if (present) {
                    if (coerced || escaped) pass = 1; /* key-forming string must be plain */
                    flags |= F_HAS_NAME;
                    row->nspan = mk_span(v->off, v->len);
                    row->nh = orc_hash64(s + v->off, v->len);
                }
            } else if (KEY("message")) {
                int present, coerced, escaped;
                if (node_to_string(v, &present, &coerced, &escaped)) {
                    err = 1;
                    break;
                }
                if (present) {
                    flags |= F_HAS_MESSAGE;
                    if (coerced) flags |= F_NAME_COERCED;
                    if (escaped) flags |= F_STR_ESCAPED;
                    row->mspan = mk_span(v->off, v->len);
                }
            } else if (KEY("source")) {
                int ord = 0, isnull;
                int rc = node_to_enum(s, v, E_SOURCE, 2, &ord, &isnull);
                if (rc == 1) err = 1;
                if (rc == 2) pass = 1;
                if (rc == 0) {
                    if (isnull)
                        flags |= F_NULL_SOURCE;
                    else
                        source = (uint32_t)ord;
                }
            } else if (KEY("level")) {
                int ord = 0, isnull;
                int rc = node_to_enum(s, v, E_LEVEL, 4, &ord, &isnull);
                if (rc == 1) err = 1;
                if (rc == 2) pass = 1;
                if (rc == 0) {
                    if (isnull)
                        flags |= F_NULL_LEVEL;
                    else
                        level = (uint32_t)ord;
                }
            }
#undef KEY
        }
        if (err) goto done;
        if (pass) {
            status = ST_PASS;
            goto done;
        }
        if (kind == KIND_MEASUREMENT) flags |= F_HAS_V0; /* primitive double always has a value */
        if (flags & F_NULL_SOURCE) source = 0;
        if (flags & F_NULL_LEVEL) level = 0;
        row->meta = mk_meta(kind, flags, level, source, has_etype, etype, rec);
        status = ST_OK;
    }
done:
    jfree_all(&p);
    return status;
}

uint64_t orc_decode_json(const uint8_t *bytes, const uint64_t *rec_off, uint32_t n, uint64_t max_rows, const orc_cols *o,
                         uint8_t *status, uint32_t *first_row) {
    uint64_t row = 0;
    for (uint32_t i = 0; i < n; i++) {
        first_row[i] = (uint32_t)row;
        uint64_t len = rec_off[i + 1] - rec_off[i];
        jrow r;
        int st = decode_one(bytes + rec_off[i], len > 0x7fffffffu ? 0x7fffffffu : (uint32_t)len, i, &r);
        status[i] = (uint8_t)st;
        if (st != ST_OK) continue;
        if (row < max_rows) {
            o->meta[row] = r.meta;
            o->token_hash[row] = r.th;
            o->name_hash[row] = r.nh;
            o->event_date[row] = r.date;
            o->v0[row] = r.v0;
            o->v1[row] = r.v1;
            o->v2[row] = r.v2;
            o->token_span[row] = r.tspan;
            o->name_span[row] = r.nspan;
            if (o->originator_span) o->originator_span[row] = r.ospan;
            if (o->message_span) o->message_span[row] = r.mspan;
            if (o->metadata_span) o->metadata_span[row] = r.mdspan;
            if (o->alternate_id_span) o->alternate_id_span[row] = r.aspan;
        }
        row++;
    }
    first_row[n] = (uint32_t)row;
    return row;
}
