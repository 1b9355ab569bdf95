"""Known-answer vectors transcribed from the reference's own tests (inputs -> expected
outputs only; paths relative to /root/reference).  Shared by the oracle pinning tests
(CPU) and the CUDA parity tests (GPU)."""

# test/test_storage.py:16-32
IND2PTR = dict(row=[2, 2, 4, 5, 5, 6], M=8, rowptr=[0, 0, 0, 2, 2, 3, 5, 6, 6])

# test/test_storage.py:35-50
SORT_ON_CONSTRUCT = dict(row=[0, 0, 1, 1], col=[1, 0, 1, 0], value=[2, 1, 4, 3],
                         out_row=[0, 0, 1, 1], out_col=[0, 1, 0, 1], out_value=[1, 2, 3, 4],
                         sizes=(2, 2))

# test/test_storage.py:53-108
CACHING = dict(row=[0, 0, 1, 1], col=[0, 1, 0, 1], rowcount=[2, 2], rowptr=[0, 2, 4],
               colcount=[2, 2], colptr=[0, 2, 4], csr2csc=[0, 2, 1, 3], csc2csr=[0, 2, 1, 3])

# test/test_storage.py:111-154
SET_VALUE = dict(row=[0, 0, 1, 1], col=[1, 0, 1, 0], value=[1, 2, 3, 4],
                 csc_layout=[1, 3, 2, 4], coo_layout=[1, 2, 3, 4])

# test/test_storage.py:157-180
STORAGE_COALESCE = dict(row=[0, 0, 0, 1, 1], col=[0, 1, 1, 0, 1], value=[1, 1, 1, 3, 4],
                        out_row=[0, 0, 1, 1], out_col=[0, 1, 0, 1], out_value=[1, 2, 3, 4])

# test/test_coalesce.py:6-34
COALESCE = dict(row=[1, 0, 1, 0, 2, 1], col=[0, 1, 1, 1, 0, 0], m=3, n=2,
                value=[[1, 2], [2, 3], [3, 4], [4, 5], [5, 6], [6, 7]],
                index=[[0, 1, 1, 2], [1, 0, 1, 0]],
                add=[[6, 8], [7, 9], [3, 4], [5, 6]],
                max=[[4, 5], [6, 7], [3, 4], [5, 6]])

# test/test_transpose.py:15-51
TRANSPOSE_MATRIX = dict(row=[1, 0, 1, 2], col=[0, 1, 1, 0], value=[1, 2, 3, 4], m=3, n=2,
                        index=[[0, 0, 1, 1], [1, 2, 0, 1]], out_value=[1, 4, 2, 3])
TRANSPOSE = dict(row=[1, 0, 1, 0, 2, 1], col=[0, 1, 1, 1, 0, 0], m=3, n=2,
                 value=[[1, 2], [2, 3], [3, 4], [4, 5], [5, 6], [6, 7]],
                 index=[[0, 0, 1, 1], [1, 2, 0, 1]],
                 out_value=[[7, 9], [5, 6], [6, 8], [3, 4]])

# test/test_sample.py:17-37
SAMPLE_ADJ = dict(row=[0, 0, 0, 1, 1, 2, 2, 2, 2, 3, 4, 4, 5, 5],
                  col=[1, 2, 3, 0, 2, 0, 1, 4, 5, 0, 2, 5, 2, 4], sizes=(6, 6),
                  subset=[2, 3, 4, 5],
                  all_n_id=[2, 3, 4, 5, 0, 1], all_row=[0, 0, 0, 0, 1, 2, 2, 3, 3],
                  all_col=[2, 3, 4, 5, 4, 0, 3, 0, 2], all_val=[7, 8, 5, 6, 9, 10, 11, 12, 13],
                  nnz_replace_k2=8, nnz_noreplace_k2=7)

# test/test_permute.py:10-21  (permute = index_select(0).index_select(1), permute.py:8-13)
PERMUTE = dict(row=[0, 0, 1, 2, 2], col=[0, 1, 0, 1, 2], value=[1, 2, 3, 4, 5],
               perm=[1, 0, 2], out_row=[0, 1, 1, 2, 2], out_col=[1, 0, 1, 0, 2],
               out_value=[3, 2, 1, 4, 5])

# test/test_tensor.py:71-88 exercises to_symmetric (a "next" row) — not pinned here.

# test/test_add.py:15-39 (sparse + sparse)
SPARSE_ADD = dict(rowA=[0, 0, 1, 2, 2], colA=[0, 2, 1, 0, 1], valueA=[1, 2, 4, 1, 3],
                  rowB=[0, 0, 1, 2, 2], colB=[1, 2, 2, 1, 2], valueB=[2, 3, 1, 2, 4],
                  row=[0, 0, 0, 1, 1, 2, 2, 2], col=[0, 1, 2, 1, 2, 0, 1, 2], value=[1, 2, 5, 4, 1, 1, 5, 4])

# test/test_mul.py:15-39 and :42-61 (sparse (.) sparse; second case has an empty result)
SPARSE_MUL = dict(rowA=[0, 0, 1, 2, 2], colA=[0, 2, 1, 0, 1], valueA=[1, 2, 4, 1, 3],
                  rowB=[0, 0, 1, 2, 2], colB=[1, 2, 2, 1, 2], valueB=[2, 3, 1, 2, 4],
                  row=[0, 2], col=[2, 1], value=[6, 6])
SPARSE_MUL_EMPTY = dict(rowA=[0], colA=[1], valueA=[1], rowB=[1], colB=[0], valueB=[2], row=[], col=[], value=[])

# test/test_tensor.py:71-88
TO_SYMMETRIC = dict(row=[0, 0, 0, 1, 1], col=[0, 1, 2, 0, 2], value=[1, 2, 3, 4, 5],
                    dense=[[2, 6, 3], [6, 0, 5], [3, 5, 0]])

# test/test_cat.py:11-66 (value-less operands with filled caches; dense view and which structures survive)
CAT = dict(row1=[0, 0, 1], col1=[0, 1, 2], row2=[0, 0, 1, 2], col2=[0, 1, 1, 0],
           dim0=[[1, 1, 0], [0, 0, 1], [1, 1, 0], [0, 1, 0], [1, 0, 0]], dim0_cached=1,
           dim1=[[1, 1, 0, 1, 1], [0, 0, 1, 0, 1], [0, 0, 0, 1, 0]], dim1_cached=2,
           diag=[[1, 1, 0, 0, 0], [0, 0, 1, 0, 0], [0, 0, 0, 1, 1], [0, 0, 0, 0, 1], [0, 0, 0, 1, 0]],
           diag_cached=5, value_width=4, value_cat_width=8, value_cat_cached=5)

# test/test_convert.py:9-30 (round trips through scipy and the framework's own sparse COO type)
CONVERT = dict(index=[[0, 0, 1, 2, 2], [0, 2, 1, 0, 1]], value=[1, 2, 4, 1, 3], N=3)

# test/test_eye.py:14-66
EYE = dict(square=dict(M=3, N=3, row=[0, 1, 2], rowptr=[0, 1, 2, 3], col=[0, 1, 2], value=[1, 1, 1]),
           wide=dict(M=3, N=4, row=[0, 1, 2], rowptr=[0, 1, 2, 3], col=[0, 1, 2], rowcount=[1, 1, 1],
                     colptr=[0, 1, 2, 3, 3], colcount=[1, 1, 1, 0], csr2csc=[0, 1, 2], csc2csr=[0, 1, 2]),
           tall=dict(M=4, N=3, row=[0, 1, 2], rowptr=[0, 1, 2, 3, 3], col=[0, 1, 2], rowcount=[1, 1, 1, 0],
                     colptr=[0, 1, 2, 3], colcount=[1, 1, 1], csr2csc=[0, 1, 2], csc2csr=[0, 1, 2]))

# test/test_overload.py:6-22 (value-less matrix, dense operand on either side of + and *)
OVERLOAD = dict(row=[0, 1, 1, 2, 2], col=[1, 0, 2, 1, 2], other=[1, 2, 3])

# test/test_storage.py:183-200 (test_sparse_reshape)
SPARSE_RESHAPE = dict(row=[0, 1, 2, 3], col=[0, 1, 2, 3],
                      steps=[((2, 8), (2, 8), [0, 0, 1, 1], [0, 5, 2, 7]),
                             ((-1, 4), (4, 4), [0, 1, 2, 3], [0, 1, 2, 3])])
