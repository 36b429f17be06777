/* _pkpack: the lock-step driver's request packer in C (CPython extension, built by pokeai_b200/build.py with gcc).
 *
 * pack_request_into(buf, offset, battle_status, request, sp_idx, mv_idx, it_idx, max_actions,
 *                   status_code, status_fnt, weather_code) -> list of simulator keys
 *
 * The same walk as pokeai_b200.protocol.pack_request_into (the reference's get_possible_actions, pokeai/ai/common.py:
 * 143-174, and StateFeatureExtractor's reads, state_feature_extractor.py:113-197, folded into one pass that writes the
 * 144-byte pk_obs record of include/pokeai_b200.h): same bytes, same keys, same exception types.  Host-side only; the
 * Python function stays as the fallback and as the specification the tests compare against. */
#define PY_SSIZE_T_CLEAN
#include <Python.h>
#include <stdint.h>
#include <string.h>

#define PK_OBS_BYTES 144
#define PK_MAX_MOVES 4
#define FLAG_SWITCH 1
#define FLAG_FORCE 2

static PyObject *s_side_statuses, *s_side_friend, *s_side_opponent, *s_weather, *s_active, *s_total_pokes,
    *s_remaining_pokes, *s_species, *s_hp_current, *s_hp_max, *s_status, *s_ranks, *s_forceSwitch, *s_trapped, *s_side,
    *s_pokemon, *s_details, *s_item, *s_moves, *s_disabled, *s_id, *s_move, *s_condition, *s_rank_keys[7];

static int intern_all(void) {
#define IN(v, t)                        \
  v = PyUnicode_InternFromString(t);    \
  if (v == NULL) return -1;
  IN(s_side_statuses, "side_statuses") IN(s_side_friend, "side_friend") IN(s_side_opponent, "side_opponent")
  IN(s_weather, "weather") IN(s_active, "active") IN(s_total_pokes, "total_pokes") IN(s_remaining_pokes, "remaining_pokes")
  IN(s_species, "species") IN(s_hp_current, "hp_current") IN(s_hp_max, "hp_max") IN(s_status, "status") IN(s_ranks, "ranks")
  IN(s_forceSwitch, "forceSwitch") IN(s_trapped, "trapped") IN(s_side, "side") IN(s_pokemon, "pokemon")
  IN(s_details, "details") IN(s_item, "item") IN(s_moves, "moves") IN(s_disabled, "disabled") IN(s_id, "id")
  IN(s_move, "move") IN(s_condition, "condition")
  static const char* rk[7] = {"atk", "def", "spa", "spd", "spe", "accuracy", "evasion"};
  for (int i = 0; i < 7; ++i) { IN(s_rank_keys[i], rk[i]) }
#undef IN
  return 0;
}

/* obj.name as a C long; -1 with an exception set on failure (values here are never negative except ranks) */
static int attr_long(PyObject* o, PyObject* name, long* out) {
  PyObject* v = PyObject_GetAttr(o, name);
  if (v == NULL) return -1;
  *out = PyLong_AsLong(v);
  Py_DECREF(v);
  return (*out == -1 && PyErr_Occurred()) ? -1 : 0;
}

/* d[key] -> long (KeyError like the Python code when the name is unknown) */
static int lookup_long(PyObject* d, PyObject* key, long* out) {
  PyObject* v = PyObject_GetItem(d, key); /* new reference; raises KeyError(key) */
  if (v == NULL) return -1;
  *out = PyLong_AsLong(v);
  Py_DECREF(v);
  return (*out == -1 && PyErr_Occurred()) ? -1 : 0;
}

/* d.get(key) truthiness for a dict-like request node; 0 when absent */
static int dict_truth(PyObject* d, PyObject* key) {
  PyObject* v = PyDict_Check(d) ? PyDict_GetItemWithError(d, key) : NULL; /* borrowed */
  if (v == NULL) {
    if (PyErr_Occurred()) return -1;
    if (!PyDict_Check(d)) {
      PyObject* w = PyObject_GetItem(d, key);
      if (w == NULL) {
        if (PyErr_ExceptionMatches(PyExc_KeyError)) {
          PyErr_Clear();
          return 0;
        }
        return -1;
      }
      int t = PyObject_IsTrue(w);
      Py_DECREF(w);
      return t;
    }
    return 0;
  }
  return PyObject_IsTrue(v);
}

static void put_u16(unsigned char* p, long v) {
  p[0] = (unsigned char)(v & 0xff);
  p[1] = (unsigned char)((v >> 8) & 0xff);
}

/* status string -> code: "" / None -> 0, known -> status_code[s], anything else -> status_fnt */
static int status_of(PyObject* st, PyObject* status_code, long status_fnt, long* out) {
  if (st == Py_None) {
    *out = 0;
    return 0;
  }
  if (PyUnicode_Check(st) && PyUnicode_GetLength(st) == 0) {
    *out = 0;
    return 0;
  }
  PyObject* v = PyDict_GetItemWithError(status_code, st); /* borrowed */
  if (v == NULL) {
    if (PyErr_Occurred()) return -1;
    *out = status_fnt;
    return 0;
  }
  *out = PyLong_AsLong(v);
  return (*out == -1 && PyErr_Occurred()) ? -1 : 0;
}

/* one side of the header: remaining / total / hp / status / ranks of side index i (0 friend, 1 opponent) */
static int pack_side(unsigned char* rec, int i, PyObject* ss, PyObject* status_code, long status_fnt, PyObject** active_out) {
  PyObject* act = PyObject_GetAttr(ss, s_active);
  if (act == NULL) return -1;
  if (act == Py_None) {
    Py_DECREF(act);
    PyErr_SetString(PyExc_AssertionError, "active is None"); /* state_feature_extractor.py:133,148,161,177 */
    return -1;
  }
  long total, remaining, hpc, hpm, stc;
  if (attr_long(ss, s_total_pokes, &total) || attr_long(ss, s_remaining_pokes, &remaining) ||
      attr_long(act, s_hp_current, &hpc) || attr_long(act, s_hp_max, &hpm)) {
    Py_DECREF(act);
    return -1;
  }
  if (total == 0 || hpm == 0) {
    Py_DECREF(act);
    PyErr_SetString(PyExc_ZeroDivisionError, "division by zero"); /* the reference divides in Python (:122,:150) */
    return -1;
  }
  PyObject* st = PyObject_GetAttr(act, s_status);
  if (st == NULL) {
    Py_DECREF(act);
    return -1;
  }
  const int rcs = status_of(st, status_code, status_fnt, &stc);
  Py_DECREF(st);
  if (rcs) {
    Py_DECREF(act);
    return -1;
  }
  PyObject* ranks = PyObject_GetAttr(act, s_ranks);
  if (ranks == NULL) {
    Py_DECREF(act);
    return -1;
  }
  for (int k = 0; k < 7; ++k) {
    long r;
    if (lookup_long(ranks, s_rank_keys[k], &r)) {
      Py_DECREF(ranks);
      Py_DECREF(act);
      return -1;
    }
    rec[14 + 7 * i + k] = (unsigned char)(signed char)r;
  }
  Py_DECREF(ranks);
  rec[0 + i] = (unsigned char)remaining;
  rec[2 + i] = (unsigned char)total;
  put_u16(rec + 4 + 2 * i, hpc);
  put_u16(rec + 8 + 2 * i, hpm);
  rec[12 + i] = (unsigned char)stc;
  *active_out = act; /* new reference handed to the caller */
  return 0;
}

/* "move 3" / "switch 5" */
#define N_CACHED_KEYS 8
static PyObject* s_key_cache[2][N_CACHED_KEYS]; /* "move 1".."move 8", "switch 1".."switch 8", made on first use */
static int add_key(PyObject* keys, const char* what, long n) {
  const int w = what[0] == 's';
  if (n >= 1 && n <= N_CACHED_KEYS) {
    if (s_key_cache[w][n - 1] == NULL) s_key_cache[w][n - 1] = PyUnicode_FromFormat("%s %ld", what, n);
    if (s_key_cache[w][n - 1] == NULL) return -1;
    return PyList_Append(keys, s_key_cache[w][n - 1]);
  }
  PyObject* k = PyUnicode_FromFormat("%s %ld", what, n);
  if (k == NULL) return -1;
  const int rc = PyList_Append(keys, k);
  Py_DECREF(k);
  return rc;
}

/* common.py:24-29: return102 -> return, frustration1 -> frustration; returns a NEW reference */
static PyObject* side_move_id(PyObject* id) {
  if (PyUnicode_Check(id)) {
    Py_ssize_t n;
    const char* s = PyUnicode_AsUTF8AndSize(id, &n);
    if (s == NULL) return NULL;
    if (n >= 6 && memcmp(s, "return", 6) == 0) return PyUnicode_FromString("return");
    if (n >= 11 && memcmp(s, "frustration", 11) == 0) return PyUnicode_FromString("frustration");
  }
  Py_INCREF(id);
  return id;
}

/* common.py:17-21: id "hiddenpower" + name "Hidden Power Dark 70" -> hiddenpowerdark; returns a NEW reference */
static PyObject* active_move_id(PyObject* id, PyObject* move) {
  if (PyUnicode_Check(id) && PyUnicode_CompareWithASCIIString(id, "hiddenpower") == 0) {
    PyObject* name = PyObject_GetItem(move, s_move);
    if (name == NULL) return NULL;
    PyObject* parts = PyUnicode_Split(name, NULL, -1); /* the Python code splits on " "; names carry single spaces */
    Py_DECREF(name);
    if (parts == NULL) return NULL;
    if (PyList_GET_SIZE(parts) < 3) {
      Py_DECREF(parts);
      PyErr_SetString(PyExc_IndexError, "list index out of range");
      return NULL;
    }
    PyObject* lower = PyObject_CallMethod(PyList_GET_ITEM(parts, 2), "lower", NULL);
    Py_DECREF(parts);
    if (lower == NULL) return NULL;
    PyObject* out = PyUnicode_FromFormat("hiddenpower%U", lower);
    Py_DECREF(lower);
    return out;
  }
  Py_INCREF(id);
  return id;
}

static PyObject* pack_request_into(PyObject* self, PyObject* args) {
  (void)self;
  PyObject *bufobj, *bs, *request, *sp_idx, *mv_idx, *it_idx, *status_code, *weather_code;
  Py_ssize_t offset;
  long max_actions, status_fnt;
  if (!PyArg_ParseTuple(args, "OnOOOOOlOlO", &bufobj, &offset, &bs, &request, &sp_idx, &mv_idx, &it_idx, &max_actions,
                        &status_code, &status_fnt, &weather_code))
    return NULL;
  if (!PyDict_Check(status_code) || !PyDict_Check(weather_code)) {
    PyErr_SetString(PyExc_TypeError, "status_code and weather_code must be dicts");
    return NULL;
  }
  unsigned char rec[PK_OBS_BYTES];
  memset(rec, 0, sizeof(rec));
  PyObject *keys = NULL, *sides = NULL, *kf = NULL, *ko = NULL, *af = NULL, *ao = NULL, *pokemon = NULL, *active = NULL,
           *act0 = NULL, *act_moves = NULL;
  int ok = 0;
  /* ---- state half (state_feature_extractor.py:113-197 reads) ---- */
  sides = PyObject_GetAttr(bs, s_side_statuses);
  kf = PyObject_GetAttr(bs, s_side_friend);
  ko = PyObject_GetAttr(bs, s_side_opponent);
  if (sides == NULL || kf == NULL || ko == NULL) goto done;
  {
    PyObject* sf = PyObject_GetItem(sides, kf);
    if (sf == NULL) goto done;
    PyObject* so = PyObject_GetItem(sides, ko);
    if (so == NULL) {
      Py_DECREF(sf);
      goto done;
    }
    const int r0 = pack_side(rec, 0, sf, status_code, status_fnt, &af);
    const int r1 = r0 ? -1 : pack_side(rec, 1, so, status_code, status_fnt, &ao);
    Py_DECREF(sf);
    Py_DECREF(so);
    if (r0 || r1) goto done;
  }
  {
    PyObject* sp = PyObject_GetAttr(ao, s_species);
    if (sp == NULL) goto done;
    long opp;
    const int rc = lookup_long(sp_idx, sp, &opp); /* KeyError like dex.get_pokedex_by_name */
    Py_DECREF(sp);
    if (rc) goto done;
    rec[28] = (unsigned char)opp;
    PyObject* w = PyObject_GetAttr(bs, s_weather);
    if (w == NULL) goto done;
    PyObject* wc = PyDict_GetItemWithError(weather_code, w); /* borrowed; absent -> 0 */
    Py_DECREF(w);
    if (wc == NULL && PyErr_Occurred()) goto done;
    rec[29] = wc ? (unsigned char)PyLong_AsLong(wc) : 0;
  }
  /* ---- actions (common.py:143-174) ---- */
  const int force_switch = dict_truth(request, s_forceSwitch);
  if (force_switch < 0) goto done;
  int trapped = 0;
  if (!force_switch) {
    active = PyObject_GetItem(request, s_active); /* KeyError when the request has no 'active' */
    if (active == NULL) goto done;
    act0 = PySequence_GetItem(active, 0);
    if (act0 == NULL) goto done;
    trapped = dict_truth(act0, s_trapped);
    if (trapped < 0) goto done;
  }
  {
    PyObject* side = PyObject_GetItem(request, s_side);
    if (side == NULL) goto done;
    pokemon = PyObject_GetItem(side, s_pokemon);
    Py_DECREF(side);
    if (pokemon == NULL) goto done;
  }
  keys = PyList_New(0);
  if (keys == NULL) goto done;
  const Py_ssize_t n_members = PySequence_Size(pokemon);
  if (n_members < 0) goto done;
  unsigned char* pos = rec + 32;
  for (Py_ssize_t pi = 0; pi < n_members; ++pi) {
    PyObject* member = PySequence_GetItem(pokemon, pi);
    if (member == NULL) goto done;
    int fail = 1;
    PyObject *details = NULL, *name = NULL, *item = NULL, *moves = NULL, *cond = NULL;
    long species, item_idx;
    details = PyObject_GetItem(member, s_details);
    if (details == NULL) goto member_done;
    {
      /* "Kangaskhan, L50, F" -> "Kangaskhan" (common.py:146-147) */
      const Py_ssize_t comma = PyUnicode_FindChar(details, ',', 0, PyUnicode_GetLength(details), 1);
      if (comma == -2) goto member_done;
      name = comma < 0 ? (Py_INCREF(details), details) : PyUnicode_Substring(details, 0, comma);
      if (name == NULL) goto member_done;
    }
    if (lookup_long(sp_idx, name, &species)) goto member_done;
    const int is_active = dict_truth(member, s_active);
    if (is_active < 0) goto member_done;
    if (is_active) {
      if (force_switch) {
        fail = 0;
        goto member_done;
      }
      item = PyObject_GetItem(member, s_item);
      if (item == NULL || lookup_long(it_idx, item, &item_idx)) goto member_done;
      if (act_moves == NULL) {
        act_moves = PyObject_GetItem(act0, s_moves);
        if (act_moves == NULL) goto member_done;
      }
      const Py_ssize_t nm = PySequence_Size(act_moves);
      if (nm < 0) goto member_done;
      for (Py_ssize_t mi = 0; mi < nm; ++mi) {
        PyObject* mv = PySequence_GetItem(act_moves, mi);
        if (mv == NULL) goto member_done;
        const int dis = dict_truth(mv, s_disabled);
        if (dis != 0) {
          Py_DECREF(mv);
          if (dis < 0) goto member_done;
          continue;
        }
        if (PyList_GET_SIZE(keys) >= max_actions) {
          Py_DECREF(mv);
          PyErr_Format(PyExc_ValueError, "more than %ld possible actions", max_actions);
          goto member_done;
        }
        PyObject* id = PyObject_GetItem(mv, s_id);
        PyObject* mid = id ? active_move_id(id, mv) : NULL;
        Py_XDECREF(id);
        Py_DECREF(mv);
        if (mid == NULL) goto member_done;
        long midx;
        const int rc = lookup_long(mv_idx, mid, &midx);
        Py_DECREF(mid);
        if (rc) goto member_done;
        pos[0] = 0;
        pos[1] = (unsigned char)species;
        pos[2] = (unsigned char)item_idx;
        pos[3] = 1;
        put_u16(pos + 4, midx);
        pos += 12;
        if (add_key(keys, "move", (long)mi + 1)) goto member_done;
      }
    } else {
      cond = PyObject_GetItem(member, s_condition);
      if (cond == NULL) goto member_done;
      {
        Py_ssize_t cl;
        const char* cs = PyUnicode_AsUTF8AndSize(cond, &cl);
        if (cs == NULL) goto member_done;
        if ((cl >= 4 && memcmp(cs + cl - 4, " fnt", 4) == 0) || trapped) {
          fail = 0;
          goto member_done;
        }
      }
      moves = PyObject_GetItem(member, s_moves);
      if (moves == NULL) goto member_done;
      const Py_ssize_t nm = PySequence_Size(moves);
      if (nm < 0) goto member_done;
      if (nm > PK_MAX_MOVES) {
        PyErr_SetString(PyExc_ValueError, "more than 4 moves on one action");
        goto member_done;
      }
      if (PyList_GET_SIZE(keys) >= max_actions) {
        PyErr_Format(PyExc_ValueError, "more than %ld possible actions", max_actions);
        goto member_done;
      }
      for (Py_ssize_t mi = 0; mi < nm; ++mi) {
        PyObject* x = PySequence_GetItem(moves, mi);
        PyObject* mid = x ? side_move_id(x) : NULL;
        Py_XDECREF(x);
        if (mid == NULL) goto member_done;
        long midx;
        const int rc = lookup_long(mv_idx, mid, &midx);
        Py_DECREF(mid);
        if (rc) goto member_done;
        put_u16(pos + 4 + 2 * mi, midx);
      }
      item = PyObject_GetItem(member, s_item);
      if (item == NULL || lookup_long(it_idx, item, &item_idx)) goto member_done;
      pos[0] = (unsigned char)(FLAG_SWITCH | (force_switch ? FLAG_FORCE : 0));
      pos[1] = (unsigned char)species;
      pos[2] = (unsigned char)item_idx;
      pos[3] = (unsigned char)nm;
      pos += 12;
      if (add_key(keys, "switch", (long)pi + 1)) goto member_done;
    }
    fail = 0;
  member_done:
    Py_XDECREF(details);
    Py_XDECREF(name);
    Py_XDECREF(item);
    Py_XDECREF(moves);
    Py_XDECREF(cond);
    Py_DECREF(member);
    if (fail) goto done;
  }
  rec[30] = (unsigned char)PyList_GET_SIZE(keys);
  {
    Py_buffer view;
    if (PyObject_GetBuffer(bufobj, &view, PyBUF_WRITABLE | PyBUF_SIMPLE)) goto done;
    if (offset < 0 || offset + PK_OBS_BYTES > view.len) {
      PyBuffer_Release(&view);
      PyErr_SetString(PyExc_ValueError, "record does not fit the buffer");
      goto done;
    }
    memcpy((unsigned char*)view.buf + offset, rec, PK_OBS_BYTES);
    PyBuffer_Release(&view);
  }
  ok = 1;
done:
  Py_XDECREF(sides);
  Py_XDECREF(kf);
  Py_XDECREF(ko);
  Py_XDECREF(af);
  Py_XDECREF(ao);
  Py_XDECREF(pokemon);
  Py_XDECREF(active);
  Py_XDECREF(act0);
  Py_XDECREF(act_moves);
  if (!ok) {
    Py_XDECREF(keys);
    return NULL;
  }
  return keys;
}

static PyMethodDef methods[] = {
    {"pack_request_into", pack_request_into, METH_VARARGS,
     "pack_request_into(buf, offset, battle_status, request, sp_idx, mv_idx, it_idx, max_actions, status_code, "
     "status_fnt, weather_code) -> simulator keys; writes the 144-byte pk_obs record at buf[offset:]"},
    {NULL, NULL, 0, NULL}};
static struct PyModuleDef moddef = {PyModuleDef_HEAD_INIT, "_pkpack", "request -> packed record (C)", -1, methods,
                                    NULL, NULL, NULL, NULL};
PyMODINIT_FUNC PyInit__pkpack(void) {
  if (intern_all()) return NULL;
  return PyModule_Create(&moddef);
}
