"""The ``metacells.extensions`` drop-in (dropin/metacells): with ``dropin/`` ahead of the installed package on sys.path,
``import metacells.extensions as xt`` serves the hot-path symbols from libmcb200.so and everything else from the stock
module; the rest of the stock package is untouched.  CPU only: names are resolved, nothing is computed on a GPU."""
import os
import shutil
import subprocess
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shim_resolves_every_path_symbol_and_falls_through(tmp_path, ref):
    stock = tmp_path / "site" / "metacells"
    (stock / "utilities").mkdir(parents=True)
    (stock / "__init__.py").write_text("STOCK_MARKER = 'stock package'\nfrom . import utilities\n")
    (stock / "utilities" / "__init__.py").write_text("from . import computation\n")
    # the name construction of utilities/computation.py:461, :879-883, :1291, :1341, :1374-1378, :1700, :278-283, :317-322
    (stock / "utilities" / "computation.py").write_text(textwrap.dedent('''
        import metacells.extensions as xt
        def path_symbols():
            names = [f"correlate_dense_{d}_t" for d in ("float32", "float64")]
            names += [f"cross_correlate_dense_{d}_t" for d in ("float32", "float64")]
            names += [f"pairs_correlate_dense_{d}_t" for d in ("float32", "float64")]
            names += ["downsample_compressed_%s_t_%s_t_%s_t" % ("float32", "int32", "float32")]
            names += ["downsample_dense_%s_t_%s_t" % ("float32", "float32"), "downsample_array_%s_t_%s_t" % ("float32", "float32")]
            names += [f"rank_matrix_{'float32'}_t", f"rank_rows_{'float32'}_t", f"collect_top_{'float32'}_t"]
            names += ["collect_pruned_%s_t_%s_t_%s_t" % ("float32", "int32", "int32")]
            names += ["collect_compressed_%s_t_%s_t_%s_t" % ("float32", "int32", "int32")]
            names += ["sort_compressed_indices_%s_t_%s_t_%s_t" % ("float32", "int32", "int32")]
            names += ["optimize_partitions", "score_partitions", "choose_seeds", "set_threads_count"]
            return {name: getattr(xt, name) for name in names}
    '''))
    have_stock_extension = ref is not None
    if have_stock_extension:
        shutil.copy(ref.__file__, stock / os.path.basename(ref.__file__))
    script = textwrap.dedent(f'''
        import sys
        sys.path[:0] = [{os.path.join(ROOT, "dropin")!r}, {ROOT!r}, {str(tmp_path / "site")!r}]
        import metacells
        assert metacells.STOCK_MARKER == "stock package"
        import metacells.extensions as xt
        from metacells.utilities import computation
        assert computation.xt is xt and xt.__file__.startswith({os.path.join(ROOT, "dropin")!r})
        symbols = computation.path_symbols()
        for name, function in symbols.items():
            assert function.__module__ in ("metacells_b200.extensions", "metacells.extensions"), (name, function.__module__)
        print("path symbols", len(symbols))
        if {have_stock_extension!r}:
            other = xt.downsample_array_float64_t_float64_t  # not on the path: the stock module's
            assert "built-in" in repr(other) or "PyCapsule" in repr(type(other)) or other.__module__ != "metacells_b200.extensions"
            import numpy as np
            values = np.arange(10, dtype="float64")
            out = np.zeros(10, dtype="float64")
            other(values, out, 20, 123456)
            assert out.sum() == 20
            xt.set_threads_count(2)
            print("fell through")
        try:
            xt.no_such_symbol_float32_t
        except AttributeError:
            print("missing raises")
    ''')
    result = subprocess.run([sys.executable, "-c", script], capture_output=True, text=True, timeout=300)
    assert result.returncode == 0, result.stdout + result.stderr
    assert "path symbols 19" in result.stdout and "missing raises" in result.stdout
    if have_stock_extension:
        assert "fell through" in result.stdout
