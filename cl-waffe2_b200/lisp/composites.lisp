;;;; composites.lisp — the nn / optimizer composites of the MLP path as single launches on CUDATensor.
;;;;
;;;; Loaded AFTER `cl-waffe2/nn` and `cl-waffe2/optimizers` (INTEGRATION.md §2: second module entry), because it specialises
;;;; things those systems define.  Nothing here changes user code: `(softmax-cross-entropy x labels)`, `(Adam p :lr ...)`,
;;;; `(SGD p :lr ...)` and `(call-optimizer! p)` keep their meaning; with CUDATensor first in `*using-backend*` they run as
;;;;   * Softmax-Cross-Entropy-Node forward  : wf2cu_softmax + ONE fused element-wise chain (instead of the two compiled
;;;;     sub-graphs of `nn/criterion.lisp:89-117`: `!softmax` = 5 nodes + 2 disguised reductions, `cross-entropy-loss` = 4 nodes)
;;;;   * Softmax-Cross-Entropy-Node backward : ONE fused chain `sce-diff-static` (`nn/criterion.lisp:81-87`: 3 nodes)
;;;;   * Adam `step-optimize`                : ONE wf2cu_adam_step per parameter (instead of three `defmodel-as` composites,
;;;;     `optimizers/impls/adam.lisp:45-116`: 9 memory passes and 4 host-made ScalarTensors per parameter per step)
;;;;   * SGD `step-optimize`                 : ONE wf2cu_sgd_step per parameter (`optimizers/impls/sgd.lisp:20-35`)
;;;; The per-element operations and their order are those of the reference's nodes (csrc/fused.cu lists every program next to
;;;; the lines it restates; csrc/misc.cu `adam_kernel`), so results equal the node-by-node path bit for bit except for the
;;;; softmax row sums (tree order instead of sequential, SURVEY §7 hard part d).
;;;; The Python host (cl-waffe2_b200/nn.py) drives the same entry points and is what runs in the tests; the matcher logic of
;;;; fusion.lisp has a Python mirror in tests/test_lisp_fusion_mirror.py.  No Common Lisp exists in the authoring image.
(in-package :cl-waffe2/backends.cuda)

;; `run-fused-program`, `float-bits` and the +p-...+ opcodes: fusion.lisp.

;; ---- Softmax-Cross-Entropy-Node (`nn/criterion.lisp:89-117`) ---------------------------------------------------------------
;; `define-op` declared the node with a `:device t` implementation that calls two compiled sub-graphs; a CUDATensor
;; implementation registered through the same `define-impl-op` wins the dispatch (`vm/nodes/defnode.lisp:215-232`).
;; The node keeps its constructor arguments (axis delta avoid-overflow) only inside those two closures, not in slots.  The
;; constructor evaluates `(determine-facet-of-nodes name *using-backend* axis delta avoid-overflow)` — which calls `:reject-p`
;; with exactly those values (`defnode.lisp:75-89,215-232,404-406`) — as an argument of the `make-instance` that follows, so
;; `:reject-p` notes them and an `initialize-instance :after` method on the node class files them under the new instance.
(defvar *sce-pending-args* nil "constructor arguments seen by the last :reject-p call (same thread, immediately before make-instance)")
(defvar *sce-args* (trivial-garbage:make-weak-hash-table :weakness :key) "node -> (axis delta avoid-overflow)")

(defun sce-reject-p (&rest constructor-args)
  (setq *sce-pending-args* (copy-list constructor-args))
  (not *cuda-available-p*))

(defmethod initialize-instance :after ((node cl-waffe2/nn::Softmax-Cross-Entropy-Node) &key &allow-other-keys)
  (when *sce-pending-args*
    (setf (gethash node *sce-args*) *sce-pending-args*
	  *sce-pending-args* nil)))

(defun sce-args (node)
  "(values axis delta avoid-overflow) of NODE; the constructor's defaults when the node was made before this file was loaded."
  (destructuring-bind (&optional (axis 1) (delta 1e-7) (avoid-overflow t)) (gethash node *sce-args*)
    (values axis (coerce delta 'single-float) avoid-overflow)))

(defvar *sce-buffers* (trivial-garbage:make-weak-hash-table :weakness :key)
  "node -> (z out): the two (rows cols) results of a forward call, reused from step to step while the shape is unchanged
(an impl-op runs at execution time, outside the compiled graph's memory pool).")

(defun sce-buffers (node x)
  (let ((cached (gethash node *sce-buffers*)))
    (if (and cached (equal (shape (first cached)) (shape x)))
	cached
	(setf (gethash node *sce-buffers*)
	      (list (make-tensor (shape x) :dtype (dtype x) :order (order x))
		    (make-tensor (shape x) :dtype (dtype x) :order (order x)))))))

(define-impl-op (cl-waffe2/nn::Softmax-Cross-Entropy-Node :device CUDATensor :reject-p #'sce-reject-p)
  :forward ((self x label)
	    (cl-waffe2/vm.nodes::with-composite-node-mode self
	      (destructuring-bind (z out) (sce-buffers self x)
		(multiple-value-bind (axis delta avoid-overflow) (sce-args self)
		  ;; z = (!softmax x :axis axis :avoid-overflow avoid-overflow): one pass
		  (with-lowered-views ((xv x) (zv z))
		    (softmax xv zv axis (if avoid-overflow 1 0)))
		  ;; out = (!mul -1 (!mul labels (!loge (!add z delta))))   (`cross-entropy-loss`, criterion.lisp:157)
		  (run-fused-program
		   `((,+p-load+ 0 0 0) (,+p-imm+ 1 0 ,(float-bits delta)) (,+p-add+ 0 0 1) (,+p-unary+ 0 0 ,+u-loge+)
		     (,+p-load+ 2 1 0) (,+p-mul+ 0 2 0) (,+p-imm+ 1 0 ,(float-bits -1.0)) (,+p-mul+ 0 1 0) (,+p-store+ 0 2 0))
		   (list z label out)))
		(with-setting-save4bw ((z z) (label label))
		  out))))
  :backward ((self dout x label)
	     ;; as `define-op` writes it (`vm/nodes/static-node.lisp:311-318`): hand over to the backward node
	     (let ((in-shape  (map 'list #'shape (list x label)))
		   (out-shape (cl-waffe2/vm.nodes::node-out-sizes self)))
	       (call (cl-waffe2/nn::Softmax-Cross-Entropy-Node-backward self in-shape out-shape) dout x label))))

(define-impl-op (cl-waffe2/nn::Softmax-Cross-Entropy-Node-backward :device CUDATensor :reject-p #'cuda-only-reject-p)
  :forward ((self dout x label)
	    (declare (ignore label))
	    (let ((fw-self (slot-value self 'cl-waffe2/vm.nodes::fw-self)))
	      (cl-waffe2/vm.nodes::with-composite-node-mode fw-self
		(with-reading-save4bw ((z z) (labels* label))
		  ;; `sce-diff-static` (criterion.lisp:81-87): dx = (dy * (z - labels)) / batch-size, batch-size = (car (last (shape z) 2))
		  (let ((dx (make-tensor (shape x) :dtype (dtype x) :order (order x)))
			(batch (coerce (car (last (shape z) 2)) 'single-float)))
		    (run-fused-program
		     `((,+p-load+ 0 0 0) (,+p-load+ 1 1 0) (,+p-sub+ 0 0 1) (,+p-load+ 2 2 0) (,+p-mul+ 0 2 0)
		       (,+p-imm+ 1 0 ,(float-bits batch)) (,+p-div+ 0 0 1) (,+p-store+ 0 3 0))
		     (list z labels* dout dx))
		    ;; gradients of (x label): labels receive none, as in the reference (`sce-diff-static` returns X.grad only)
		    (values dx nil)))))))

;; ---- optimizers: `step-optimize` (`optimizers/defoptimizer.lisp:46`) -------------------------------------------------------
;; The reference keeps one optimizer object per parameter; `(call-optimizer! p)` calls `step-optimize` on it
;; (`vm/generic-tensor/tensor.lisp:1182-1197`).  An :around method keeps every slot and the step counter of the reference's
;; classes and replaces only the arithmetic when the parameter lives on the device.
(defun dense-cuda-p (&rest tensors)
  (every #'(lambda (x) (and (typep x 'CUDATensor) (not (scalar-p x)))) tensors))

(defun flat-count (tensor) (apply #'* (mapcar #'observe (shape tensor))))

(defmethod cl-waffe2/optimizers:step-optimize :around ((optimizer cl-waffe2/optimizers:Adam))
  (let* ((param (cl-waffe2/optimizers:read-parameter optimizer))
	 (g     (grad param))
	 (m     (cl-waffe2/optimizers::adam-m optimizer))
	 (v     (cl-waffe2/optimizers::adam-v optimizer)))
    (if (and *cuda-available-p* g (dense-cuda-p param g m v))
	;; m <- m + (g - m)(1 - b1) ; v <- v + (1 - b2)(g g - v) ; p <- p - (m lr_t) / (eps + sqrt v), as Adam-Step-M / -V /
	;; -Param write it (adam.lisp:45-67); lr_t from the reference's own `adam-step-lr` (it also bumps N, adam.lisp:81-94)
	(let ((lr-t (coerce (cl-waffe2/optimizers::adam-step-lr optimizer) 'single-float)))
	  (adam-step (device-vec-ptr (tensor-vec param)) (tensor-initial-offset param)
		     (device-vec-ptr (tensor-vec g)) (tensor-initial-offset g)
		     (device-vec-ptr (tensor-vec m)) (tensor-initial-offset m)
		     (device-vec-ptr (tensor-vec v)) (tensor-initial-offset v)
		     (flat-count param) lr-t (null-pointer)
		     (coerce (cl-waffe2/optimizers::beta1-of optimizer) 'single-float)
		     (coerce (cl-waffe2/optimizers::beta2-of optimizer) 'single-float)
		     (coerce (cl-waffe2/optimizers::eps-of optimizer) 'single-float)
		     1.0)
	  param)
	(call-next-method))))

(defmethod cl-waffe2/optimizers:step-optimize :around ((optimizer cl-waffe2/optimizers:SGD))
  (let* ((param (cl-waffe2/optimizers:read-parameter optimizer))
	 (g     (grad param)))
    (if (and *cuda-available-p* g (dense-cuda-p param g))
	;; p <- p - (g * lr)   (`SGD-Compute-Form`, sgd.lisp:20-26; the reference's A*=scal also leaves g * lr in g: the next
	;; backward overwrites it, nothing reads it in between)
	(progn
	  (sgd-step (device-vec-ptr (tensor-vec param)) (tensor-initial-offset param)
		    (device-vec-ptr (tensor-vec g)) (tensor-initial-offset g)
		    (flat-count param) (coerce (cl-waffe2/optimizers::sgd-lr optimizer) 'single-float))
	  param)
	(call-next-method))))
