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

;; The example only runs the model forward (`gpt2-inference`, model.lisp:216-239) and its weights are loaded once by the
;; model constructors (model.lisp:46-160): declaring the weight matrices constant lets MatmulNode keep their tensor-core
;; operand planes between calls (wf2cu_gemm_const_operand) instead of re-splitting 48 matrices per forward.
;; Call once after `(GPT2)` when the tensors live on CUDATensor; a later upload into a weight re-splits it.
(defun declare-constant-weights (model &key (constant t))
  (flet ((declare-one (tensor)
	   (when (typep tensor 'cl-waffe2/backends.cuda:CUDATensor)
	     (cl-waffe2/backends.cuda::gemm-const-operand
	      (cl-waffe2/backends.cuda::device-vec-ptr (tensor-vec tensor))
	      (tensor-initial-offset tensor)
	      (apply #'* (shape tensor))
	      (if constant 1 0)))))
    (declare-one (slot-value model 'wte))
    (dolist (layer (slot-value model 'layers))
      (dolist (slot '(attn-attn-w attn-proj-w mlp-fc-w mlp-proj-w))
	(declare-one (slot-value layer slot))))
    model))
