ssh test@127.0.0.1 "kill -9 5770"
ssh test@127.0.0.1 "kill -9 5771"
ssh test@127.0.0.1 "kill -9 5772"
kill -9 5044
