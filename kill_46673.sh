ssh test@127.0.0.1 "kill -9 5234"
ssh test@127.0.0.1 "kill -9 5235"
ssh test@127.0.0.1 "kill -9 5236"
kill -9 5044
