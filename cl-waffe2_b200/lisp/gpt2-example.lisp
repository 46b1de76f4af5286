;;;; gpt2-example.lisp — CUDATensor implementation of the GPT-2 example's own node.
;;;;
;;;; Loaded with the example (`examples/gpt-2/`, after its utils.lisp), not with the backend: the node belongs to the
;;;; :gpt-2-example package.  `GPT2PositionalEmbedding` (`examples/gpt-2/utils.lisp:65-112`) has only a LispTensor
;;;; implementation that walks the sentence on the host and calls `waffe2-sadd` on host pointers twice per token
;;;; (`expand-embedding-form`, utils.lisp:19-62); with this file and CUDATensor first in `*using-backend*` the same node is
;;;; one gather launch:  ctx-out[b, p, :] = wte[round(ctx[b, p]), :] + wpe[p, :]   (token ids are floats, as in the example).
(in-package :gpt-2-example)

(define-impl-op (GPT2PositionalEmbedding :device cl-waffe2/backends.cuda:CUDATensor
					 :reject-p #'cl-waffe2/backends.cuda::cuda-only-reject-p)
  :forward ((self ctx wte wpe ctx-out)
	    (declare (ignore self))
	    (assert (and (eql (order ctx) :column) (eql (order wpe) :column) (eql (order wte) :column))
		    nil
		    "GPT2PositionalEmbedding: Orders must be :column (C Order), not a :row (Fortran Order)")
	    (flet ((ptr (tensor)
		     ;; dense operands: the storage pointer advanced by the view's first element
		     (cffi:inc-pointer (cl-waffe2/backends.cuda::device-vec-ptr (tensor-vec tensor))
				       (* 4 (tensor-initial-offset tensor)))))
	      (cl-waffe2/backends.cuda::embedding-gather
	       (ptr ctx)
	       (cl-waffe2/vm:maybe-observe-axis (first (shape ctx)))
	       (cl-waffe2/vm:maybe-observe-axis (second (shape ctx)))
	       (ptr wte) (ptr wpe)
	       (cl-waffe2/vm:maybe-observe-axis (second (shape wte)))
	       (ptr ctx-out)))
	    ctx-out))
