ssh test@127.0.0.1 "kill -9 9599"
ssh test@127.0.0.1 "kill -9 9600"
ssh test@127.0.0.1 "kill -9 9601"
kill -9 8873
