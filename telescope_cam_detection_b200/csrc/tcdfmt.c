/* _tcdfmt -- CPython helper of YOLOXDetector._format_model_output_to_detections.
 *
 * Builds the reference's detection dicts (src/yolox_detector.py:222-282) from a result table in C: at ~100
 * detections per frame the dict construction was the largest host cost of the coordinator path (~1.2 us per dict in
 * a Python comprehension, the whole budget of a frame at 8 000 frames/s; with 8 ranks sharing one host it bounded
 * the end-to-end scaling).  Same arithmetic as the reference, operation for operation:
 *   ratio   = orig / network size                  (Python float division, double)
 *   x1..y2  = float(np.float32(row[i]) * ratio)    (NumPy 2 weak scalar: the ratio is narrowed to float32, the
 *                                                   product is float32, then widened)
 *   conf    = float(np.float32(obj) * np.float32(cls))
 *   area    = int((x2 - x1) * (y2 - y1))           (double arithmetic, truncation toward zero)
 * Pinned by tests/golden/format_ref.json (produced by the reference's own method).  Host logic only -- nothing
 * here touches the GPU path.
 */
#define PY_SSIZE_T_CLEAN
#include <Python.h>
#include <stdint.h>

static PyObject *k_class_id, *k_class_name, *k_confidence, *k_bbox, *k_x1, *k_y1, *k_x2, *k_y2, *k_area;

static int wildlife(long c) { return c == 0 || c == 14 || c == 15 || c == 16 || c == 21; }

/* format_rows(rows: C-contiguous float32 buffer [n,7], orig_h: float, orig_w: float, in_h: int, in_w: int,
 *             wildlife_only: bool, names: tuple[str]) -> list[dict] */
static PyObject* format_rows(PyObject* self, PyObject* args) {
  Py_buffer buf;
  double orig_h, orig_w;
  long in_h, in_w;
  int wl;
  PyObject* names;
  if (!PyArg_ParseTuple(args, "y*ddllpO!", &buf, &orig_h, &orig_w, &in_h, &in_w, &wl, &PyTuple_Type, &names)) return NULL;
  if (buf.len % (7 * sizeof(float)) != 0) {
    PyBuffer_Release(&buf);
    PyErr_SetString(PyExc_ValueError, "rows must be a float32 [n,7] buffer");
    return NULL;
  }
  const Py_ssize_t n = buf.len / (Py_ssize_t)(7 * sizeof(float));
  const float* rows = (const float*)buf.buf;
  const float rw = (float)(orig_w / (double)in_w), rh = (float)(orig_h / (double)in_h);
  const Py_ssize_t n_names = PyTuple_GET_SIZE(names);
  PyObject* out = PyList_New(0);
  if (!out) { PyBuffer_Release(&buf); return NULL; }
  for (Py_ssize_t i = 0; i < n; ++i) {
    const float* r = rows + 7 * i;
    const long cid = (long)r[6];
    if (wl && !wildlife(cid)) continue;
    const float fx1 = r[0] * rw, fy1 = r[1] * rh, fx2 = r[2] * rw, fy2 = r[3] * rh, fc = r[4] * r[5];
    const double x1 = (double)fx1, y1 = (double)fy1, x2 = (double)fx2, y2 = (double)fy2;
    const double a = (x2 - x1) * (y2 - y1);
    PyObject *det = PyDict_New(), *bbox = PyDict_New(), *v = NULL;
    if (!det || !bbox) goto fail_item;
#define SET(d, k, expr) do { v = (expr); if (!v || PyDict_SetItem(d, k, v) < 0) goto fail_item; Py_DECREF(v); v = NULL; } while (0)
    SET(det, k_class_id, PyLong_FromLong(cid));
    if (cid >= 0 && cid < n_names) {
      if (PyDict_SetItem(det, k_class_name, PyTuple_GET_ITEM(names, cid)) < 0) goto fail_item;
    } else {
      SET(det, k_class_name, PyUnicode_FromFormat("class_%ld", cid));
    }
    SET(det, k_confidence, PyFloat_FromDouble((double)fc));
    SET(bbox, k_x1, PyFloat_FromDouble(x1));
    SET(bbox, k_y1, PyFloat_FromDouble(y1));
    SET(bbox, k_x2, PyFloat_FromDouble(x2));
    SET(bbox, k_y2, PyFloat_FromDouble(y2));
    /* int(double): truncation toward zero; PyLong_FromDouble does exactly that (and raises on inf / nan like int()) */
    SET(bbox, k_area, PyLong_FromDouble(a));
#undef SET
    if (PyDict_SetItem(det, k_bbox, bbox) < 0) goto fail_item;
    Py_DECREF(bbox);
    bbox = NULL;
    if (PyList_Append(out, det) < 0) goto fail_item;
    Py_DECREF(det);
    continue;
  fail_item:
    Py_XDECREF(v);
    Py_XDECREF(det);
    Py_XDECREF(bbox);
    Py_DECREF(out);
    PyBuffer_Release(&buf);
    return NULL;
  }
  PyBuffer_Release(&buf);
  return out;
}

static PyMethodDef methods[] = {
    {"format_rows", format_rows, METH_VARARGS, "postprocess rows [n,7] float32 -> reference detection dicts"},
    {NULL, NULL, 0, NULL}};

static struct PyModuleDef moddef = {PyModuleDef_HEAD_INIT, "_tcdfmt", "detection-dict builder (host logic)", -1, methods};

PyMODINIT_FUNC PyInit__tcdfmt(void) {
  k_class_id = PyUnicode_InternFromString("class_id");
  k_class_name = PyUnicode_InternFromString("class_name");
  k_confidence = PyUnicode_InternFromString("confidence");
  k_bbox = PyUnicode_InternFromString("bbox");
  k_x1 = PyUnicode_InternFromString("x1");
  k_y1 = PyUnicode_InternFromString("y1");
  k_x2 = PyUnicode_InternFromString("x2");
  k_y2 = PyUnicode_InternFromString("y2");
  k_area = PyUnicode_InternFromString("area");
  return PyModule_Create(&moddef);
}
